// b200vq -- streaming entropy regulariser (reference models/bottleneck.py:12-33 applied to -d, :300-303).
//
// The reference materialises A = -d / T_e (N x K fp32), softmax(A), log_softmax(A) and their autograd
// buffers.  Here A[i][k] = (2 z_i.e_k - |z_i|^2 - |e_k|^2) / T_e is recomputed tile by tile in exact fp32
// and never stored.  One kernel template serves five passes; a thread owns one vector (a latent row or a
// code row, held in registers) and the other side streams through shared memory 64 vectors at a time:
//
//   forward   ROWSTATS   rows own, codes stream : online (max, sum e^a, sum e^a a) -> lse_i, ent_i = sum_k p log p
//             COLSUM     codes own, rows stream : psum_k = sum_i exp(a_ik - lse_i)            (= N * mean prob)
//             finalize   one CTA over K         : avg entropy, w_k = d(avg entropy)/d(pbar_k), the 3 scalars
//   backward  ROWDOT     rows own, codes stream : c_i = sum_k p_ik w_k
//             GRAD_ROWS  rows own, codes stream : dz_i  = (2/T) sum_k g_ik (e_k - z_i)
//             GRAD_CODES codes own, rows stream : de_k  = (2/T) sum_i g_ik (z_i - e_k)
//   with g_ik = p_ik / N * ( -cS (log p_ik - ent_i) + cA (w_k - c_i) ),  cS = g_loss + g_sample,
//   cA = g_avg - g_loss  (loss = sample_entropy - avg_entropy, reference :32).
#pragma once
#include "vq_common.cuh"

namespace b200vq {

enum : int { ENT_ROWSTATS = 0, ENT_COLSUM = 1, ENT_ROWDOT = 2, ENT_GRAD_ROWS = 3, ENT_GRAD_CODES = 4 };
constexpr int kEntTile = 64;
constexpr int kEntThreads = 128;
constexpr int kEntMaxOwn = 2;        // owner vectors per thread: 2 on big problems, 1 when CTAs are scarce
constexpr int kEntChunk = 8;
constexpr float kEntNegBig = -1e30f;
constexpr float kEntPadSq = 1e25f;   // |v|^2 of a padding vector: its affinity underflows every exp()
constexpr float kEntSkip = 1e-14f;   // chunks whose largest probability is below this add nothing in fp32
constexpr float kEntEps = 1e-5f;     // reference :30

struct EntropyArgs {
    const float* own;      // (n_own, D) owner vectors
    int64_t n_own;
    const float* str;      // (n_str, D) streamed vectors
    int64_t n_str;
    int64_t str_per_split; // streamed vectors per blockIdx.y (multiple of kEntTile)
    float t_inv;           // 1 / T_e
    float inv_n;           // 1 / number of latent rows
    const float* own_a;    // per-owner scalars (meaning depends on the pass, see the table above)
    const float* own_b;
    const float* own_c;
    const float* str_a;    // per-streamed-vector scalars
    const float* str_b;
    const float* str_c;
    const float* g3;       // device [g_loss, g_sample, g_avg] (gradient passes)
    float* out_a;
    float* out_b;
    float* out_vec;        // (n_own, D)
    double* out_sum;
};

// R owner vectors per thread: every streamed vector read from shared memory (one broadcast LDS.128 per four
// elements) feeds R dot products -- and, in the gradient passes, the rank-1 accumulations as well, from the
// same registers -- so the FMA pipe, not the shared-memory wavefront rate, is the limit.
// gridDim.y > 1 splits the streamed side; partial results are then merged with atomics (sums) or, for ROWSTATS,
// written as (max, sum e^a, sum e^a a) triples to out_vec[split][owner][3] for entropy_merge_kernel.
template <int D, int PASS, int R>
__global__ void __launch_bounds__(kEntThreads) entropy_pass_kernel(const EntropyArgs p) {
    constexpr int D4 = D / 4;
    constexpr bool kGrad = PASS == ENT_GRAD_ROWS || PASS == ENT_GRAD_CODES;
    __shared__ __align__(16) float s_vec[kEntTile * D];
    __shared__ float s_sq[kEntTile], s_a[kEntTile], s_b[kEntTile], s_c[kEntTile];
    __shared__ double s_red[kEntThreads / 32];

    const int tid = threadIdx.x;
    int64_t own_i[R];
    bool live[R];
    float o[R][D], o2[R], oa[R], ob[R], oc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        own_i[r] = (static_cast<int64_t>(blockIdx.x) * R + r) * kEntThreads + tid;
        live[r] = own_i[r] < p.n_own;
        const float4* op = reinterpret_cast<const float4*>(p.own + (live[r] ? own_i[r] : 0) * D);
#pragma unroll
        for (int q = 0; q < D4; ++q) {
            const float4 v = live[r] ? __ldg(op + q) : make_float4(0.f, 0.f, 0.f, 0.f);
            o[r][4 * q + 0] = v.x; o[r][4 * q + 1] = v.y; o[r][4 * q + 2] = v.z; o[r][4 * q + 3] = v.w;
        }
        o2[r] = 0.f;
#pragma unroll
        for (int d = 0; d < D; ++d) o2[r] = fmaf(o[r][d], o[r][d], o2[r]);
        oa[r] = (live[r] && p.own_a) ? __ldg(p.own_a + own_i[r]) : 0.f;
        ob[r] = (live[r] && p.own_b) ? __ldg(p.own_b + own_i[r]) : 0.f;
        oc[r] = (live[r] && p.own_c) ? __ldg(p.own_c + own_i[r]) : 0.f;
    }
    float cS = 0.f, cA = 0.f;
    if (kGrad) {
        const float gl = __ldg(p.g3 + 0), gs = __ldg(p.g3 + 1), ga = __ldg(p.g3 + 2);
        cS = gl + gs;
        cA = ga - gl;
    }

    float m[R], l[R], s[R];          // ROWSTATS
    float acc0[R];                    // COLSUM / ROWDOT
    float acc[R][kGrad ? D : 1];      // gradient passes
    float accg[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        m[r] = kEntNegBig; l[r] = 0.f; s[r] = 0.f; acc0[r] = 0.f; accg[r] = 0.f;
#pragma unroll
        for (int d = 0; d < (kGrad ? D : 1); ++d) acc[r][d] = 0.f;
    }

    const int64_t s_begin = static_cast<int64_t>(blockIdx.y) * p.str_per_split;
    const int64_t s_end = min(p.n_str, s_begin + p.str_per_split);
    for (int64_t t0 = s_begin; t0 < s_end; t0 += kEntTile) {
        const int64_t left = s_end - t0;
        const int cnt = left < kEntTile ? static_cast<int>(left) : kEntTile;
        __syncthreads();
        {
            const float4* src = reinterpret_cast<const float4*>(p.str + t0 * D);
            float4* dst = reinterpret_cast<float4*>(s_vec);
            const int n4 = cnt * D4;
            for (int q = tid; q < kEntTile * D4; q += kEntThreads)
                dst[q] = q < n4 ? __ldg(src + q) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (tid < kEntTile) {
                const bool ok = tid < cnt;
                s_a[tid] = (ok && p.str_a) ? __ldg(p.str_a + t0 + tid) : 0.f;
                s_b[tid] = (ok && p.str_b) ? __ldg(p.str_b + t0 + tid) : 0.f;
                s_c[tid] = (ok && p.str_c) ? __ldg(p.str_c + t0 + tid) : 0.f;
            }
        }
        __syncthreads();
        if (tid < kEntTile) {
            float sq = 0.f;
#pragma unroll
            for (int d = 0; d < D; ++d) { const float v = s_vec[tid * D + d]; sq = fmaf(v, v, sq); }
            s_sq[tid] = tid < cnt ? sq : kEntPadSq;
        }
        __syncthreads();

        if constexpr (!kGrad) {
#pragma unroll 1
            for (int j0 = 0; j0 < cnt; j0 += kEntChunk) {
                float a[R][kEntChunk];
#pragma unroll
                for (int jj = 0; jj < kEntChunk; ++jj) {
                    const float4* v4 = reinterpret_cast<const float4*>(s_vec + (j0 + jj) * D);
                    float dot[R];
#pragma unroll
                    for (int r = 0; r < R; ++r) dot[r] = 0.f;
#pragma unroll
                    for (int q = 0; q < D4; ++q) {
                        const float4 v = v4[q];
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            dot[r] = fmaf(o[r][4 * q + 0], v.x, dot[r]);
                            dot[r] = fmaf(o[r][4 * q + 1], v.y, dot[r]);
                            dot[r] = fmaf(o[r][4 * q + 2], v.z, dot[r]);
                            dot[r] = fmaf(o[r][4 * q + 3], v.w, dot[r]);
                        }
                    }
                    const float sq = s_sq[j0 + jj];
#pragma unroll
                    for (int r = 0; r < R; ++r) a[r][jj] = (2.f * dot[r] - (o2[r] + sq)) * p.t_inv;
                }
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if (PASS == ENT_ROWSTATS) {
                        float cm = a[r][0];
#pragma unroll
                        for (int jj = 1; jj < kEntChunk; ++jj) cm = fmaxf(cm, a[r][jj]);
                        const float mn = fmaxf(m[r], cm);
                        const float sc = __expf(m[r] - mn);
                        l[r] *= sc;
                        s[r] *= sc;
#pragma unroll
                        for (int jj = 0; jj < kEntChunk; ++jj) {
                            const float e = __expf(a[r][jj] - mn);
                            l[r] += e;
                            s[r] = fmaf(e, a[r][jj], s[r]);
                        }
                        m[r] = mn;
                    } else if (PASS == ENT_COLSUM) {
#pragma unroll
                        for (int jj = 0; jj < kEntChunk; ++jj) acc0[r] += __expf(a[r][jj] - s_a[j0 + jj]);
                    } else {
#pragma unroll
                        for (int jj = 0; jj < kEntChunk; ++jj)
                            acc0[r] = fmaf(__expf(a[r][jj] - oa[r]), s_a[j0 + jj], acc0[r]);
                    }
                }
            }
        } else {
            // gradient passes: one streamed vector at a time, held in registers for the dot products AND the
            // rank-1 accumulation; the accumulation is skipped when no lane of the warp sees a probability
            // above kEntSkip (most (row, code) pairs at the reference's T_e = 0.01)
#pragma unroll 2
            for (int j = 0; j < cnt; ++j) {
                float v[D];
                const float4* v4 = reinterpret_cast<const float4*>(s_vec + j * D);
#pragma unroll
                for (int q = 0; q < D4; ++q) {
                    const float4 t = v4[q];
                    v[4 * q + 0] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
                }
                const float sq = s_sq[j], sa = s_a[j], sb = s_b[j], sc = s_c[j];
                float g[R];
                float pm = 0.f;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    float dot = 0.f;
#pragma unroll
                    for (int d = 0; d < D; ++d) dot = fmaf(o[r][d], v[d], dot);
                    const float a = (2.f * dot - (o2[r] + sq)) * p.t_inv;
                    // the latent row's (lse, ent, c) and the code's w: owner or streamed side by pass
                    const float rl = PASS == ENT_GRAD_ROWS ? oa[r] : sa;
                    const float re = PASS == ENT_GRAD_ROWS ? ob[r] : sb;
                    const float rc = PASS == ENT_GRAD_ROWS ? oc[r] : sc;
                    const float w = PASS == ENT_GRAD_ROWS ? sa : oa[r];
                    const float logp = a - rl;
                    const float pr = __expf(logp);
                    pm = fmaxf(pm, pr);
                    g[r] = pr * fmaf(-cS, logp - re, cA * (w - rc));
                }
                if (__any_sync(0xffffffffu, pm > kEntSkip)) {
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        accg[r] += g[r];
#pragma unroll
                        for (int d = 0; d < D; ++d) acc[r][d] = fmaf(g[r], v[d], acc[r][d]);
                    }
                }
            }
        }
    }

    if constexpr (PASS == ENT_ROWSTATS) {
        if (gridDim.y > 1) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (live[r]) {
                    float* dst = p.out_vec + (static_cast<int64_t>(blockIdx.y) * p.n_own + own_i[r]) * 3;
                    dst[0] = m[r]; dst[1] = l[r]; dst[2] = s[r];
                }
            }
            return;
        }
        float entv = 0.f;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (live[r]) {
                const float lse = m[r] + __logf(l[r]);
                const float e = s[r] / l[r] - lse;
                p.out_a[own_i[r]] = lse;
                p.out_b[own_i[r]] = e;
                entv += e;
            }
        }
        double red = warp_sum_d(static_cast<double>(entv));
        if ((tid & 31) == 0) s_red[tid >> 5] = red;
        __syncthreads();
        if (tid == 0) {
            double t = 0.0;
            for (int w = 0; w < kEntThreads / 32; ++w) t += s_red[w];
            atomicAdd(p.out_sum, t);
        }
    } else if constexpr (PASS == ENT_COLSUM) {
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (live[r] && s_begin < s_end) atomicAdd(p.out_a + own_i[r], acc0[r]);
    } else if constexpr (PASS == ENT_ROWDOT) {
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (live[r]) {
                if (gridDim.y > 1) { if (s_begin < s_end) atomicAdd(p.out_a + own_i[r], acc0[r]); }
                else p.out_a[own_i[r]] = acc0[r];
            }
    } else {
        const float scale = 2.f * p.t_inv * p.inv_n;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (!(live[r] && s_begin < s_end)) continue;
            float* dst = p.out_vec + own_i[r] * D;
            if (PASS == ENT_GRAD_ROWS && gridDim.y == 1) {
#pragma unroll
                for (int q = 0; q < D4; ++q) {
                    float4 t;
                    t.x = scale * fmaf(-o[r][4 * q + 0], accg[r], acc[r][4 * q + 0]);
                    t.y = scale * fmaf(-o[r][4 * q + 1], accg[r], acc[r][4 * q + 1]);
                    t.z = scale * fmaf(-o[r][4 * q + 2], accg[r], acc[r][4 * q + 2]);
                    t.w = scale * fmaf(-o[r][4 * q + 3], accg[r], acc[r][4 * q + 3]);
                    reinterpret_cast<float4*>(dst)[q] = t;
                }
            } else {
#pragma unroll
                for (int d = 0; d < D; ++d) atomicAdd(dst + d, scale * fmaf(-o[r][d], accg[r], acc[r][d]));
            }
        }
    }
}

// ROWSTATS with a split code stream: merge the per-split (max, sum e^a, sum e^a a) triples of a row.
__global__ void __launch_bounds__(kEntThreads) entropy_merge_kernel(const float* __restrict__ part, int splits, int64_t N,
                                                                    float* __restrict__ lse_out, float* __restrict__ ent_out,
                                                                    double* __restrict__ ent_sum) {
    __shared__ double s_red[kEntThreads / 32];
    const int64_t i = static_cast<int64_t>(blockIdx.x) * kEntThreads + threadIdx.x;
    float entv = 0.f;
    if (i < N) {
        float m = kEntNegBig;
        for (int sp = 0; sp < splits; ++sp) m = fmaxf(m, part[(static_cast<int64_t>(sp) * N + i) * 3]);
        float l = 0.f, s = 0.f;
        for (int sp = 0; sp < splits; ++sp) {
            const float* t = part + (static_cast<int64_t>(sp) * N + i) * 3;
            const float sc = __expf(t[0] - m);
            l = fmaf(t[1], sc, l);
            s = fmaf(t[2], sc, s);
        }
        const float lse = m + __logf(l);
        entv = s / l - lse;
        lse_out[i] = lse;
        ent_out[i] = entv;
    }
    double red = warp_sum_d(static_cast<double>(entv));
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = red;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < kEntThreads / 32; ++w) t += s_red[w];
        atomicAdd(ent_sum, t);
    }
}

// avg entropy, its derivative per code and the three scalars.  One CTA.
__global__ void __launch_bounds__(1024) entropy_finalize_kernel(const float* __restrict__ psum, int64_t K, float inv_n,
                                                                const double* __restrict__ ent_sum,
                                                                float* __restrict__ w, float* __restrict__ out3) {
    __shared__ double s_red[32];
    double part = 0.0;
    for (int64_t k = threadIdx.x; k < K; k += blockDim.x) {
        const float pb = psum[k] * inv_n;
        const float lg = logf(pb + kEntEps);
        part += static_cast<double>(pb * lg);
        w[k] = -(lg + pb / (pb + kEntEps));
    }
    part = warp_sum_d(part);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = part;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < static_cast<int>(blockDim.x >> 5); ++i) t += s_red[i];
        const float avg = static_cast<float>(-t);
        const float sample = static_cast<float>(-(*ent_sum) * static_cast<double>(inv_n));
        out3[0] = sample - avg;
        out3[1] = sample;
        out3[2] = avg;
    }
}

}  // namespace b200vq

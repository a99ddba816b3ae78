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

template <int D, int PASS>
__global__ void __launch_bounds__(kEntThreads) entropy_pass_kernel(const EntropyArgs p) {
    constexpr int D4 = D / 4;
    constexpr bool kGrad = PASS == ENT_GRAD_ROWS || PASS == ENT_GRAD_CODES;
    __shared__ __align__(16) float s_vec[kEntTile * D];
    __shared__ float s_sq[kEntTile], s_a[kEntTile], s_b[kEntTile], s_c[kEntTile];
    __shared__ double s_red[kEntThreads / 32];

    const int tid = threadIdx.x;
    const int64_t i = static_cast<int64_t>(blockIdx.x) * kEntThreads + tid;
    const bool live = i < p.n_own;

    float o[D];
    float o2 = 0.f;
    {
        const float4* op = reinterpret_cast<const float4*>(p.own + (live ? i : 0) * D);
#pragma unroll
        for (int q = 0; q < D4; ++q) {
            const float4 v = live ? __ldg(op + q) : make_float4(0.f, 0.f, 0.f, 0.f);
            o[4 * q + 0] = v.x; o[4 * q + 1] = v.y; o[4 * q + 2] = v.z; o[4 * q + 3] = v.w;
        }
#pragma unroll
        for (int d = 0; d < D; ++d) o2 = fmaf(o[d], o[d], o2);
    }
    const float oa = (live && p.own_a) ? __ldg(p.own_a + i) : 0.f;
    const float ob = (live && p.own_b) ? __ldg(p.own_b + i) : 0.f;
    const float oc = (live && p.own_c) ? __ldg(p.own_c + i) : 0.f;
    float cS = 0.f, cA = 0.f;
    if (kGrad) {
        const float gl = __ldg(p.g3 + 0), gs = __ldg(p.g3 + 1), ga = __ldg(p.g3 + 2);
        cS = gl + gs;
        cA = ga - gl;
    }

    float m = kEntNegBig, l = 0.f, s = 0.f;   // ROWSTATS
    float acc0 = 0.f;                          // COLSUM / ROWDOT
    float acc[kGrad ? D : 1];
    float accg = 0.f;
#pragma unroll
    for (int d = 0; d < (kGrad ? D : 1); ++d) acc[d] = 0.f;

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

#pragma unroll 1
        for (int j0 = 0; j0 < cnt; j0 += kEntChunk) {
            float a[kEntChunk];
#pragma unroll
            for (int jj = 0; jj < kEntChunk; ++jj) {
                const float4* v4 = reinterpret_cast<const float4*>(s_vec + (j0 + jj) * D);
                float dot = 0.f;
#pragma unroll
                for (int q = 0; q < D4; ++q) {
                    const float4 v = v4[q];
                    dot = fmaf(o[4 * q + 0], v.x, dot);
                    dot = fmaf(o[4 * q + 1], v.y, dot);
                    dot = fmaf(o[4 * q + 2], v.z, dot);
                    dot = fmaf(o[4 * q + 3], v.w, dot);
                }
                a[jj] = (2.f * dot - (o2 + s_sq[j0 + jj])) * p.t_inv;
            }
            if (PASS == ENT_ROWSTATS) {
                float cm = a[0];
#pragma unroll
                for (int jj = 1; jj < kEntChunk; ++jj) cm = fmaxf(cm, a[jj]);
                const float mn = fmaxf(m, cm);
                const float sc = __expf(m - mn);
                l *= sc;
                s *= sc;
#pragma unroll
                for (int jj = 0; jj < kEntChunk; ++jj) {
                    const float e = __expf(a[jj] - mn);
                    l += e;
                    s = fmaf(e, a[jj], s);
                }
                m = mn;
            } else if (PASS == ENT_COLSUM) {
#pragma unroll
                for (int jj = 0; jj < kEntChunk; ++jj) acc0 += __expf(a[jj] - s_a[j0 + jj]);
            } else if (PASS == ENT_ROWDOT) {
#pragma unroll
                for (int jj = 0; jj < kEntChunk; ++jj) acc0 = fmaf(__expf(a[jj] - oa), s_a[j0 + jj], acc0);
            } else {
                float g[kEntChunk];
                float pm = 0.f;
#pragma unroll
                for (int jj = 0; jj < kEntChunk; ++jj) {
                    // the latent row's (lse, ent, c) and the code's w: owner or streamed side by pass
                    const float rl = PASS == ENT_GRAD_ROWS ? oa : s_a[j0 + jj];
                    const float re = PASS == ENT_GRAD_ROWS ? ob : s_b[j0 + jj];
                    const float rc = PASS == ENT_GRAD_ROWS ? oc : s_c[j0 + jj];
                    const float w = PASS == ENT_GRAD_ROWS ? s_a[j0 + jj] : oa;
                    const float logp = a[jj] - rl;
                    const float pr = __expf(logp);
                    pm = fmaxf(pm, pr);
                    g[jj] = pr * fmaf(-cS, logp - re, cA * (w - rc));
                }
                if (pm > kEntSkip) {
#pragma unroll
                    for (int jj = 0; jj < kEntChunk; ++jj) {
                        const float4* v4 = reinterpret_cast<const float4*>(s_vec + (j0 + jj) * D);
                        accg += g[jj];
#pragma unroll
                        for (int q = 0; q < D4; ++q) {
                            const float4 v = v4[q];
                            acc[4 * q + 0] = fmaf(g[jj], v.x, acc[4 * q + 0]);
                            acc[4 * q + 1] = fmaf(g[jj], v.y, acc[4 * q + 1]);
                            acc[4 * q + 2] = fmaf(g[jj], v.z, acc[4 * q + 2]);
                            acc[4 * q + 3] = fmaf(g[jj], v.w, acc[4 * q + 3]);
                        }
                    }
                }
            }
        }
    }

    if (PASS == ENT_ROWSTATS) {
        float entv = 0.f;
        if (live) {
            const float lse = m + __logf(l);
            entv = s / l - lse;
            p.out_a[i] = lse;
            p.out_b[i] = entv;
        }
        double r = warp_sum_d(static_cast<double>(entv));
        if ((tid & 31) == 0) s_red[tid >> 5] = r;
        __syncthreads();
        if (tid == 0) {
            double t = 0.0;
            for (int w = 0; w < kEntThreads / 32; ++w) t += s_red[w];
            atomicAdd(p.out_sum, t);
        }
    } else if (PASS == ENT_COLSUM) {
        if (live && s_begin < s_end) atomicAdd(p.out_a + i, acc0);
    } else if (PASS == ENT_ROWDOT) {
        if (live) p.out_a[i] = acc0;
    } else {
        if (live && s_begin < s_end) {
            const float scale = 2.f * p.t_inv * p.inv_n;
            float* dst = p.out_vec + i * D;
            if (PASS == ENT_GRAD_ROWS) {
#pragma unroll
                for (int q = 0; q < D4; ++q) {
                    float4 v;
                    v.x = scale * fmaf(-o[4 * q + 0], accg, acc[4 * q + 0]);
                    v.y = scale * fmaf(-o[4 * q + 1], accg, acc[4 * q + 1]);
                    v.z = scale * fmaf(-o[4 * q + 2], accg, acc[4 * q + 2]);
                    v.w = scale * fmaf(-o[4 * q + 3], accg, acc[4 * q + 3]);
                    reinterpret_cast<float4*>(dst)[q] = v;
                }
            } else {
#pragma unroll
                for (int d = 0; d < D; ++d) atomicAdd(dst + d, scale * fmaf(-o[d], accg, acc[d]));
            }
        }
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

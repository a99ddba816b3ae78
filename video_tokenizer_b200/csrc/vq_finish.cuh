// b200vq -- gather / straight-through / loss / histogram, backward, decode, statistics.
#pragma once
#include "vq_scores.cuh"

namespace b200vq {

// =======================================================================================
// Finish: keys -> indices, gather, STE output, squared-error sum, usage histograms.
// (bottleneck.py:292-298, :304-310; trainer bincounts larp_tokenizer_trainer.py:337-343)
// One thread per row; block = 128.  scratch[0] = double sum, scratch[1] = block ticket.
// =======================================================================================
template <int D, typename T>
__global__ void __launch_bounds__(128) finish_kernel(const T* __restrict__ z, int64_t z_row_stride, int64_t N,
                                                     const float* __restrict__ emb, int64_t K,
                                                     const unsigned long long* __restrict__ keys,
        const float* __restrict__ q_rows, int l2norm,
                                                     int64_t n_first, float* __restrict__ znorm_out,
                                                     float* __restrict__ ste_out, int64_t* __restrict__ idx_out,
                                                     float* __restrict__ zlen_out, int32_t* __restrict__ hist,
                                                     int32_t* __restrict__ hist_first, double* __restrict__ scratch,
                                                     int finalize, int64_t N_total, float beta, float w_cb,
                                                     float* __restrict__ losses) {
    const int64_t row = static_cast<int64_t>(blockIdx.x) * 128 + threadIdx.x;
    float sq = 0.f;
    if (row < N) {
        float zn[D];
        float z2, zlen;
        load_normalised_row<D, T>(z + row * z_row_stride, l2norm, zn, z2, zlen);
        uint32_t code = key_code(keys[row]);
        if (code >= K) code = 0;  // unreachable for valid keys; keeps the gather in bounds
        const float* q = q_rows ? q_rows + row * D : emb + static_cast<int64_t>(code) * D;
        float4* zo = reinterpret_cast<float4*>(znorm_out + row * D);   // dense (N, D) fp32, D % 4 == 0: 16-byte accesses
        float4* so = reinterpret_cast<float4*>(ste_out + row * D);
        const float4* q4 = reinterpret_cast<const float4*>(q);
#pragma unroll
        for (int d = 0; d < D; d += 4) {
            const float4 qv = __ldg(q4 + d / 4);
            const float qd[4] = {qv.x, qv.y, qv.z, qv.w};
            float st[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float diff = __fsub_rn(qd[e], zn[d + e]);
                st[e] = __fadd_rn(zn[d + e], diff);  // z + (q - z).detach()   (:307)
                sq = fmaf(diff, diff, sq);
            }
            zo[d / 4] = make_float4(zn[d], zn[d + 1], zn[d + 2], zn[d + 3]);
            so[d / 4] = make_float4(st[0], st[1], st[2], st[3]);
        }
        idx_out[row] = static_cast<int64_t>(code);
        zlen_out[row] = zlen;
        if (hist) atomicAdd(hist + code, 1);
        if (hist_first && row < n_first) atomicAdd(hist_first + code, 1);
    }
    // block reduction of the squared error, one double atomic per block
    __shared__ float s_part[4];
    sq = warp_sum(sq);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = sq;
    __syncthreads();
    if (threadIdx.x == 0) {
        const double tot = static_cast<double>(s_part[0]) + s_part[1] + s_part[2] + s_part[3];
        atomicAdd(scratch, tot);
        if (finalize) {
            __threadfence();
            unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + 1);
            const unsigned int t = atomicAdd(ticket, 1u);
            if (t == gridDim.x - 1) {
                __threadfence();
                const double s = *reinterpret_cast<volatile double*>(scratch);
                const float l = static_cast<float>(s / (static_cast<double>(N_total) * D));
                losses[1] = l;                                        // loss_commit   (:295,:297)
                losses[2] = l;                                        // loss_codebook (:296,:298)
                losses[0] = __fadd_rn(__fmul_rn(beta, l), __fmul_rn(w_cb, l));  // loss_q (:304), entropy weight 0
            }
        }
    }
}

// four consecutive elements of a dense row (16-byte aligned for fp32, 8-byte for the half types)
__device__ __forceinline__ void store4(float* p, float a, float b, float c, float d) {
    *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
__device__ __forceinline__ void store4(__nv_bfloat16* p, float a, float b, float c, float d) {
    const __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
    *reinterpret_cast<uint2*>(p) = make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
}
__device__ __forceinline__ void store4(__half* p, float a, float b, float c, float d) {
    const __half2 lo = __floats2half2_rn(a, b), hi = __floats2half2_rn(c, d);
    *reinterpret_cast<uint2*>(p) = make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
}

// =======================================================================================
// Backward.  One thread per row.  coef from device scalars so no host sync is needed.
// =======================================================================================
template <int D, typename GT>
__global__ void __launch_bounds__(128) backward_kernel(const float* __restrict__ g_ste, const float* __restrict__ g_zn,
                                                       const float* __restrict__ g_loss,
                                                       const float* __restrict__ znorm, const float* __restrict__ zlen,
                                                       const int64_t* __restrict__ idx, const float* __restrict__ emb,
                                                       const float* __restrict__ q_rows, int64_t k_offset, int64_t K_local,
                                                       int64_t N, int l2norm, int64_t N_total, float beta, float w_cb,
                                                       GT* __restrict__ grad_z, float* __restrict__ grad_emb) {
    const int64_t row = static_cast<int64_t>(blockIdx.x) * 128 + threadIdx.x;
    if (row >= N) return;
    float gq = 0.f, gc = 0.f, gb = 0.f;
    if (g_loss) { gq = __ldg(g_loss); gc = __ldg(g_loss + 1); gb = __ldg(g_loss + 2); }
    const float c = 2.0f / (static_cast<float>(N_total) * D);
    const float a = (gq * beta + gc) * c;   // commitment: d/dz  of mean((q.detach()-z)^2)
    const float b = (gq * w_cb + gb) * c;   // codebook:   d/dq  of mean((q-z.detach())^2)
    const int64_t code = idx[row] - k_offset;
    const bool owned = code >= 0 && code < K_local;
    const float* q = q_rows ? q_rows + row * D : emb + (owned ? code : 0) * D;
    float zn[D], dzh[D];
    float proj = 0.f;
    const float4* zn4 = reinterpret_cast<const float4*>(znorm + row * D);   // dense (N, D) fp32, D % 4 == 0
    const float4* q4 = reinterpret_cast<const float4*>(q);
    const float4* gs4 = reinterpret_cast<const float4*>(g_ste ? g_ste + row * D : nullptr);
    const float4* gz4 = reinterpret_cast<const float4*>(g_zn ? g_zn + row * D : nullptr);
#pragma unroll
    for (int d = 0; d < D; d += 4) {
        const float4 zv = zn4[d / 4], qv = __ldg(q4 + d / 4);
        float4 gv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (g_ste) gv = gs4[d / 4];
        if (g_zn) { const float4 t = gz4[d / 4]; gv.x += t.x; gv.y += t.y; gv.z += t.z; gv.w += t.w; }
        const float zz[4] = {zv.x, zv.y, zv.z, zv.w}, qq[4] = {qv.x, qv.y, qv.z, qv.w}, gg[4] = {gv.x, gv.y, gv.z, gv.w};
        float sc[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            zn[d + e] = zz[e];
            const float g = a * (zz[e] - qq[e]) + gg[e];
            dzh[d + e] = g;
            proj = fmaf(zz[e], g, proj);
            sc[e] = b * (qq[e] - zz[e]);
        }
        // codebook-loss gradient scatter-add: one 16-byte vector reduction per four elements (red.global.add.v4.f32)
        if (b != 0.f && owned) atomicAdd(reinterpret_cast<float4*>(grad_emb + code * D + d), make_float4(sc[0], sc[1], sc[2], sc[3]));
    }
    GT* out = grad_z + row * D;
    if (l2norm) {
        const float len = zlen[row];
        if (len > kNormEps) {
            const float inv = 1.0f / len;
#pragma unroll
            for (int d = 0; d < D; ++d) dzh[d] = (dzh[d] - zn[d] * proj) * inv;
        } else {
#pragma unroll
            for (int d = 0; d < D; ++d) dzh[d] = dzh[d] * (1.0f / kNormEps);
        }
    }
#pragma unroll
    for (int d = 0; d < D; d += 4) store4(out + d, dzh[d], dzh[d + 1], dzh[d + 2], dzh[d + 3]);
}

// grad_weight = J_normalise^T (grad_emb + g_emb).  One warp per code row.
__global__ void __launch_bounds__(256) codebook_backward_kernel(const float* __restrict__ grad_emb,
                                                                const float* __restrict__ g_emb,
                                                                const float* __restrict__ emb,
                                                                const float* __restrict__ wnorm, int64_t K, int D,
                                                                int l2norm, float* __restrict__ grad_weight) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    if (row >= K) return;
    float proj = 0.f;
    for (int d = lane; d < D; d += 32) {
        float g = grad_emb[row * D + d];
        if (g_emb) g += g_emb[row * D + d];
        proj = fmaf(emb[row * D + d], g, proj);
    }
    proj = warp_sum(proj);
    const float wn = wnorm[row];
    for (int d = lane; d < D; d += 32) {
        float g = grad_emb[row * D + d];
        if (g_emb) g += g_emb[row * D + d];
        float o = g;
        if (l2norm) o = wn > kNormEps ? (g - emb[row * D + d] * proj) / wn : g * (1.0f / kNormEps);
        grad_weight[row * D + d] = o;
    }
}

// get_codebook_entry (bottleneck.py:327-341).  One warp per index.
__global__ void __launch_bounds__(256) decode_kernel(const int64_t* __restrict__ idx, int64_t n,
                                                     const float* __restrict__ w, int64_t K, int D, int l2norm,
                                                     float* __restrict__ out, int* __restrict__ bad_flag) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    if (i >= n) return;
    const int64_t code = idx[i];
    if (code < 0 || code >= K) {
        for (int d = lane; d < D; d += 32) out[i * D + d] = __int_as_float(0x7fc00000);
        if (bad_flag && lane == 0) atomicExch(bad_flag, 1);
        return;
    }
    const float* wr = w + code * D;
    float ss = 0.f;
    for (int d = lane; d < D; d += 32) { const float v = __ldg(wr + d); ss = fmaf(v, v, ss); }
    ss = warp_sum(ss);
    const float den = l2norm ? fmaxf(sqrtf(ss), kNormEps) : 1.0f;
    for (int d = lane; d < D; d += 32) out[i * D + d] = l2norm ? __fdiv_rn(__ldg(wr + d), den) : __ldg(wr + d);
}

// Trainer usage statistics (larp_tokenizer_trainer.py:335-352).  Single block of 1024.
__global__ void __launch_bounds__(1024) usage_stats_kernel(const int32_t* __restrict__ hf, const int32_t* __restrict__ hb,
                                                           int64_t K, float* __restrict__ out4) {
    // two block reductions (counts, then the two entropies), each over a small vector so that the barriers are shared
    __shared__ double s_red[3][32];
    __shared__ double s_bcast[3];
    auto block_sum3 = [&](double (&v)[3]) {
#pragma unroll
        for (int i = 0; i < 3; ++i) v[i] = warp_sum_d(v[i]);
        __syncthreads();   // s_red / s_bcast of the previous reduction are no longer read
        if ((threadIdx.x & 31) == 0) {
#pragma unroll
            for (int i = 0; i < 3; ++i) s_red[i][threadIdx.x >> 5] = v[i];
        }
        __syncthreads();
        if (threadIdx.x < 96) {
            const double t = warp_sum_d(s_red[threadIdx.x >> 5][threadIdx.x & 31]);
            if ((threadIdx.x & 31) == 0) s_bcast[threadIdx.x >> 5] = t;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 3; ++i) v[i] = s_bcast[i];
    };
    // counts are exact in int32 (<= N < 2^31 per bin, K bins per thread-stride); the sums go through double
    int nzf_i = 0, nzb_i = 0;
    long long tot_i = 0;
    for (int64_t k = threadIdx.x; k < K; k += blockDim.x) {
        const int cf = hf[k], cb = hb[k];
        nzf_i += cf > 0; nzb_i += cb > 0; tot_i += cf;
    }
    double c3[3] = {static_cast<double>(nzf_i), static_cast<double>(nzb_i), static_cast<double>(tot_i)};
    block_sum3(c3);
    const double nzf = c3[0], nzb = c3[1], tot = c3[2];
    double kl = 0, ent = 0;
    if (tot > 0) {
        const float logK = logf(static_cast<float>(K));
        for (int64_t k = threadIdx.x; k < K; k += blockDim.x) {
            const int cf = hf[k];
            // avg_probs = one_hot.float().mean(0) -> fp32 count/n ; log(avg + 1e-10) in fp32
            const float avg = static_cast<float>(cf) / static_cast<float>(tot);
            ent -= static_cast<double>(avg * logf(avg + 1e-10f));
            // Categorical KL of the reference is fp32 arithmetic (utils/statistics.py:16-19); fp64 log() is ~40x slower here
            if (cf > 0) kl += static_cast<double>(avg * (logf(avg) + logK));
        }
    }
    double e3[3] = {kl, ent, 0.0};
    block_sum3(e3);
    if (threadIdx.x == 0) {
        out4[0] = static_cast<float>(nzf / static_cast<double>(K));
        out4[1] = static_cast<float>(nzb / static_cast<double>(K));
        out4[2] = static_cast<float>(e3[0]);
        out4[3] = static_cast<float>(exp(e3[1]));
    }
}

__global__ void __launch_bounds__(256) gather_owned_kernel(const unsigned long long* __restrict__ keys, int64_t N,
                                                           const float* __restrict__ emb_local, int64_t K_local,
                                                           int64_t k_offset, int D, float* __restrict__ out) {
    const int64_t t = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= N * D) return;
    const int64_t row = t / D;
    const int d = static_cast<int>(t - row * D);
    const int64_t code = static_cast<int64_t>(key_code(keys[row])) - k_offset;
    out[t] = (code >= 0 && code < K_local) ? emb_local[code * D + d] : 0.f;
}

__global__ void keys_reset_kernel(unsigned long long* keys, int64_t N) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < N) keys[i] = 0ull;
}

}  // namespace b200vq

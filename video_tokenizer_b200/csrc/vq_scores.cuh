// b200vq -- exact-fp32 CUDA-core kernels of the quantization bottleneck (sm_100a).
//
// These kernels are the bit-faithful path: every score is an fp32 FMA chain combined
// exactly as the reference combines it (models/bottleneck.py:264-290), so deterministic
// indices match the CPU reference outside the 1e-6 top-2-gap window.  They also serve
// as the exact rescoring stage of the tensor-core path (vq_tc.cu).
//
// The N x K score matrix is never materialised: one thread owns one token row (its
// normalised z lives in registers), code tiles stream through shared memory, and each
// row carries a running (best value, best code) pair that is merged into a 64-bit key.
#pragma once
#include "vq_common.cuh"

namespace b200vq {

// =======================================================================================
// Codebook preparation: get_emb (bottleneck.py:252-259).  One warp per code row.
// =======================================================================================
static __global__ void __launch_bounds__(256) prepare_codebook_kernel(const float* __restrict__ w, int64_t K, int D,
                                                               int l2norm, float* __restrict__ emb,
                                                               float* __restrict__ e2, float* __restrict__ wnorm) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    if (row >= K) return;
    const float* wr = w + row * D;
    float ss = 0.f;
    for (int d = lane; d < D; d += 32) { const float v = wr[d]; ss = fmaf(v, v, ss); }
    ss = warp_sum(ss);
    const float nrm = sqrtf(ss);
    const float den = l2norm ? fmaxf(nrm, kNormEps) : 1.0f;
    float s2 = 0.f;
    for (int d = lane; d < D; d += 32) {
        const float v = l2norm ? __fdiv_rn(wr[d], den) : wr[d];
        emb[row * D + d] = v;
        s2 = fmaf(v, v, s2);
    }
    s2 = warp_sum(s2);
    if (lane == 0) { e2[row] = s2; wnorm[row] = nrm; }
}

// =======================================================================================
// Row loading + normalisation shared by the scoring and finishing kernels, so that both
// see bit-identical normalised rows.
// =======================================================================================
template <int D, typename T>
__device__ __forceinline__ void load_normalised_row(const T* __restrict__ zr, int l2norm, float (&zn)[D],
                                                    float& z2, float& zlen) {
    float ss = 0.f;
#pragma unroll
    for (int d = 0; d < D; ++d) { zn[d] = to_f32(zr[d]); ss = fmaf(zn[d], zn[d], ss); }
    zlen = sqrtf(ss);
    if (l2norm) {
        const float den = fmaxf(zlen, kNormEps);
        float s2 = 0.f;
#pragma unroll
        for (int d = 0; d < D; ++d) { zn[d] = __fdiv_rn(zn[d], den); s2 = fmaf(zn[d], zn[d], s2); }
        z2 = s2;
    } else {
        z2 = ss;
    }
}

// value to MAXIMISE for code k given dot product s
template <int MODE>
__device__ __forceinline__ float rank_value(float s, float z2, float e2k, float t_inv, float g) {
    if (MODE == B200VQ_MODE_ARGMIN_DIST) {
        // d = (|z|^2 + |e|^2) - 2 s, each step rounded once like the reference's fp32 ops
        const float t = __fadd_rn(z2, e2k);
        return -__fmaf_rn(-2.0f, s, t);
    } else if (MODE == B200VQ_MODE_ARGMAX_COS) {
        return s;
    } else {
        return __fmaf_rn(s, t_inv, g);
    }
}

template <int D> struct TileCodes { static constexpr int value = D <= 32 ? 256 : (D <= 64 ? 128 : 64); };

// =======================================================================================
// Scoring pass.  grid = (row tiles of 128, code splits); block = 128 threads.
// =======================================================================================
template <int D, typename T, int MODE>
__global__ void __launch_bounds__(128) scores_kernel(const T* __restrict__ z, int64_t z_row_stride, int64_t N,
                                                     const float* __restrict__ emb, const float* __restrict__ e2,
                                                     int64_t K_local, int64_t k_offset, int64_t codes_per_split,
                                                     int l2norm, float t_inv, uint64_t seed, uint64_t offset,
                                                     int64_t row_offset, unsigned long long* __restrict__ keys) {
    constexpr int TK = TileCodes<D>::value;
    __shared__ __align__(16) float s_e[TK * D];
    __shared__ float s_e2[TK];

    const int64_t row = static_cast<int64_t>(blockIdx.x) * 128 + threadIdx.x;
    const bool row_ok = row < N;
    float zn[D];
    float z2 = 0.f, zlen = 0.f;
    if (row_ok) {
        load_normalised_row<D, T>(z + row * z_row_stride, l2norm, zn, z2, zlen);
    } else {
#pragma unroll
        for (int d = 0; d < D; ++d) zn[d] = 0.f;
    }

    const int64_t k_begin = static_cast<int64_t>(blockIdx.y) * codes_per_split;
    const int64_t k_end = min(K_local, k_begin + codes_per_split);
    float best_v = -INFINITY;
    uint32_t best_k = static_cast<uint32_t>(k_offset + k_begin);
    const uint32_t grow = static_cast<uint32_t>(row_offset + row);

    for (int64_t kt = k_begin; kt < k_end; kt += TK) {
        const int cnt = static_cast<int>(min(static_cast<int64_t>(TK), k_end - kt));
        __syncthreads();
        // cooperative, coalesced tile load (float4 when the slice base is 16B aligned)
        {
            const float* src = emb + kt * D;
            const int nflt = cnt * D;
            if ((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (nflt & 3) == 0) {
                const float4* s4 = reinterpret_cast<const float4*>(src);
                float4* d4 = reinterpret_cast<float4*>(s_e);
                for (int i = threadIdx.x; i < (nflt >> 2); i += 128) d4[i] = __ldg(s4 + i);
            } else {
                for (int i = threadIdx.x; i < nflt; i += 128) s_e[i] = __ldg(src + i);
            }
            for (int i = cnt * D + threadIdx.x; i < TK * D; i += 128) s_e[i] = 0.f;
            for (int i = threadIdx.x; i < TK; i += 128) s_e2[i] = i < cnt ? __ldg(e2 + kt + i) : 0.f;
        }
        __syncthreads();

        constexpr int STEP = MODE == B200VQ_MODE_SAMPLE ? 8 : 4;
        for (int kk = 0; kk < cnt; kk += STEP) {
            float acc[STEP];
#pragma unroll
            for (int j = 0; j < STEP; ++j) acc[j] = 0.f;
            if (D % 4 == 0) {
#pragma unroll
                for (int d = 0; d < D; d += 4) {
#pragma unroll
                    for (int j = 0; j < STEP; ++j) {
                        const float4 e = *reinterpret_cast<const float4*>(&s_e[(kk + j) * D + d]);
                        acc[j] = fmaf(zn[d + 0], e.x, acc[j]);
                        acc[j] = fmaf(zn[d + 1], e.y, acc[j]);
                        acc[j] = fmaf(zn[d + 2], e.z, acc[j]);
                        acc[j] = fmaf(zn[d + 3], e.w, acc[j]);
                    }
                }
            } else {
#pragma unroll
                for (int d = 0; d < D; ++d) {
#pragma unroll
                    for (int j = 0; j < STEP; ++j) acc[j] = fmaf(zn[d], s_e[(kk + j) * D + d], acc[j]);
                }
            }
            float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            if (MODE == B200VQ_MODE_SAMPLE) {
                // k_offset, kt and kk are multiples of 8: one noise group per step
                const uint32_t group = static_cast<uint32_t>(static_cast<uint64_t>(k_offset + kt + kk) >> 3);
                const Philox4 hw = group_head_words(group >> 3, grow, seed, offset);
                group_noise(group, grow, seed, offset, head_hi16(hw, group), g);
            }
#pragma unroll
            for (int j = 0; j < STEP; ++j) {
                const float v = rank_value<MODE>(acc[j], z2, s_e2[kk + j], t_inv, g[j]);
                if (kk + j < cnt && v > best_v) {
                    best_v = v;
                    best_k = static_cast<uint32_t>(k_offset + kt + kk + j);
                }
            }
        }
    }
    if (row_ok && k_begin < k_end) atomicMax(keys + row, make_key(best_v, best_k));
}

// Any-D fallback (D <= kMaxDim, not one of the specialised sizes): z row kept in shared
// memory, same arithmetic order.  Slow; correctness path for unusual bottleneck_dim.
template <typename T, int MODE>
__global__ void __launch_bounds__(64) scores_generic_kernel(const T* __restrict__ z, int64_t z_row_stride, int64_t N,
                                                            int D, const float* __restrict__ emb,
                                                            const float* __restrict__ e2, int64_t K_local,
                                                            int64_t k_offset, int64_t codes_per_split, int l2norm,
                                                            float t_inv, uint64_t seed, uint64_t offset,
                                                            int64_t row_offset, unsigned long long* __restrict__ keys) {
    extern __shared__ float s_z[];  // [blockDim.x][D+1]
    const int64_t row = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (row >= N) return;
    float* zn = s_z + threadIdx.x * (D + 1);
    const T* zr = z + row * z_row_stride;
    float ss = 0.f;
    for (int d = 0; d < D; ++d) { zn[d] = to_f32(zr[d]); ss = fmaf(zn[d], zn[d], ss); }
    float z2 = ss;
    if (l2norm) {
        const float den = fmaxf(sqrtf(ss), kNormEps);
        float s2 = 0.f;
        for (int d = 0; d < D; ++d) { zn[d] = __fdiv_rn(zn[d], den); s2 = fmaf(zn[d], zn[d], s2); }
        z2 = s2;
    }
    const int64_t k_begin = static_cast<int64_t>(blockIdx.y) * codes_per_split;
    const int64_t k_end = min(K_local, k_begin + codes_per_split);
    float best_v = -INFINITY;
    uint32_t best_k = static_cast<uint32_t>(k_offset + k_begin);
    const uint32_t grow = static_cast<uint32_t>(row_offset + row);
    float gn[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int64_t k = k_begin; k < k_end; ++k) {
        const float* e = emb + k * D;
        float acc = 0.f;
        for (int d = 0; d < D; ++d) acc = fmaf(zn[d], __ldg(e + d), acc);
        float g = 0.f;
        if (MODE == B200VQ_MODE_SAMPLE) {
            const uint64_t kg = static_cast<uint64_t>(k_offset + k);
            if ((kg & 7) == 0 || k == k_begin) {
                const uint32_t group = static_cast<uint32_t>(kg >> 3);
                const Philox4 hw = group_head_words(group >> 3, grow, seed, offset);
                group_noise(group, grow, seed, offset, head_hi16(hw, group), gn);
            }
            g = gn[kg & 7];
        }
        const float v = rank_value<MODE>(acc, z2, __ldg(e2 + k), t_inv, g);
        if (v > best_v) { best_v = v; best_k = static_cast<uint32_t>(k_offset + k); }
    }
    if (k_begin < k_end) atomicMax(keys + row, make_key(best_v, best_k));
}

}  // namespace b200vq

// b200vq -- any-D (runtime D <= kMaxDim) variants of finish / backward.  Correctness path
// for bottleneck_dim values that have no register-resident specialisation.
#pragma once
#include "vq_finish.cuh"

namespace b200vq {

template <typename T>
__global__ void __launch_bounds__(128) finish_generic_kernel(const T* __restrict__ z, int64_t z_row_stride, int64_t N,
                                                             int D, const float* __restrict__ emb, int64_t K,
                                                             const unsigned long long* __restrict__ keys,
        const float* __restrict__ q_rows, int l2norm,
                                                             int64_t n_first, float* __restrict__ znorm_out,
                                                             float* __restrict__ ste_out, int64_t* __restrict__ idx_out,
                                                             float* __restrict__ zlen_out, int32_t* __restrict__ hist,
                                                             int32_t* __restrict__ hist_first,
                                                             double* __restrict__ scratch, int finalize,
                                                             int64_t N_total, float beta, float w_cb,
                                                             float* __restrict__ losses) {
    const int64_t row = static_cast<int64_t>(blockIdx.x) * 128 + threadIdx.x;
    float sq = 0.f;
    if (row < N) {
        const T* zr = z + row * z_row_stride;
        float ss = 0.f;
        for (int d = 0; d < D; ++d) { const float v = to_f32(zr[d]); ss = fmaf(v, v, ss); }
        const float zlen = sqrtf(ss);
        const float den = l2norm ? fmaxf(zlen, kNormEps) : 1.0f;
        uint32_t code = key_code(keys[row]);
        if (code >= K) code = 0;
        const float* q = q_rows ? q_rows + row * D : emb + static_cast<int64_t>(code) * D;
        for (int d = 0; d < D; ++d) {
            const float v = to_f32(zr[d]);
            const float zn = l2norm ? __fdiv_rn(v, den) : v;
            const float qd = __ldg(q + d);
            const float diff = __fsub_rn(qd, zn);
            znorm_out[row * D + d] = zn;
            ste_out[row * D + d] = __fadd_rn(zn, diff);
            sq = fmaf(diff, diff, sq);
        }
        idx_out[row] = static_cast<int64_t>(code);
        zlen_out[row] = zlen;
        if (hist) atomicAdd(hist + code, 1);
        if (hist_first && row < n_first) atomicAdd(hist_first + code, 1);
    }
    __shared__ float s_part[4];
    sq = warp_sum(sq);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = sq;
    __syncthreads();
    if (threadIdx.x == 0) {
        atomicAdd(scratch, static_cast<double>(s_part[0]) + s_part[1] + s_part[2] + s_part[3]);
        if (finalize) {
            __threadfence();
            const unsigned int t = atomicAdd(reinterpret_cast<unsigned int*>(scratch + 1), 1u);
            if (t == gridDim.x - 1) {
                __threadfence();
                const double s = *reinterpret_cast<volatile double*>(scratch);
                const float l = static_cast<float>(s / (static_cast<double>(N_total) * D));
                losses[1] = l;
                losses[2] = l;
                losses[0] = __fadd_rn(__fmul_rn(beta, l), __fmul_rn(w_cb, l));
            }
        }
    }
}

template <typename GT>
__global__ void __launch_bounds__(128) backward_generic_kernel(const float* __restrict__ g_ste,
                                                               const float* __restrict__ g_zn,
                                                               const float* __restrict__ g_loss,
                                                               const float* __restrict__ znorm,
                                                               const float* __restrict__ zlen,
                                                               const int64_t* __restrict__ idx,
                                                               const float* __restrict__ emb,
                                                               const float* __restrict__ q_rows, int64_t k_offset,
                                                               int64_t K_local, int64_t N, int D,
                                                               int l2norm, int64_t N_total, float beta, float w_cb,
                                                               GT* __restrict__ grad_z, float* __restrict__ grad_emb) {
    const int64_t row = static_cast<int64_t>(blockIdx.x) * 128 + threadIdx.x;
    if (row >= N) return;
    float gq = 0.f, gc = 0.f, gb = 0.f;
    if (g_loss) { gq = __ldg(g_loss); gc = __ldg(g_loss + 1); gb = __ldg(g_loss + 2); }
    const float c = 2.0f / (static_cast<float>(N_total) * D);
    const float a = (gq * beta + gc) * c;
    const float b = (gq * w_cb + gb) * c;
    const int64_t code = idx[row] - k_offset;
    const bool owned = code >= 0 && code < K_local;
    const float* q = q_rows ? q_rows + row * D : emb + (owned ? code : 0) * D;
    float proj = 0.f;
    for (int d = 0; d < D; ++d) {
        const float zn = znorm[row * D + d];
        const float qd = __ldg(q + d);
        float g = a * (zn - qd);
        if (g_ste) g += g_ste[row * D + d];
        if (g_zn) g += g_zn[row * D + d];
        proj = fmaf(zn, g, proj);
        if (b != 0.f && owned) atomicAdd(grad_emb + code * D + d, b * (qd - zn));
    }
    const float len = l2norm ? zlen[row] : 1.f;
    for (int d = 0; d < D; ++d) {
        const float zn = znorm[row * D + d];
        float g = a * (zn - __ldg(q + d));
        if (g_ste) g += g_ste[row * D + d];
        if (g_zn) g += g_zn[row * D + d];
        float o = g;
        if (l2norm) o = len > kNormEps ? (g - zn * proj) * (1.0f / len) : g * (1.0f / kNormEps);
        grad_z[row * D + d] = from_f32<GT>(o);
    }
}

}  // namespace b200vq

// Shared device helpers for the b200vq kernels (sm_100a).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "b200vq.h"

namespace b200vq {

constexpr float kNormEps = 1e-12f;  // F.normalize eps (reference bottleneck.py:255,267)
constexpr int kMaxDim = 256;

// ---- element loads --------------------------------------------------------------------
template <int DT> struct ElemT;
template <> struct ElemT<B200VQ_DT_F32> { using type = float; };
template <> struct ElemT<B200VQ_DT_BF16> { using type = __nv_bfloat16; };
template <> struct ElemT<B200VQ_DT_F16> { using type = __half; };

__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_f32(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

// ---- selection keys -------------------------------------------------------------------
// Monotone map float -> uint32 (a < b  <=>  ord(a) < ord(b)); -0 is canonicalised first.
__device__ __forceinline__ uint32_t orderable(float v) {
    v = v + 0.0f;  // -0 -> +0 so that equal scores compare equal
    uint32_t b = __float_as_uint(v);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
// unsigned max over keys == highest value, then lowest code index
__device__ __forceinline__ unsigned long long make_key(float value, uint32_t code) {
    return (static_cast<unsigned long long>(orderable(value)) << 32) | static_cast<uint32_t>(~code);
}
__device__ __forceinline__ uint32_t key_code(unsigned long long key) { return ~static_cast<uint32_t>(key); }

// ---- Philox4x32-10 (Salmon et al., SC'11) ---------------------------------------------
struct Philox4 { uint32_t x, y, z, w; };
__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += W0; k1 += W1;
    }
    return {c0, c1, c2, c3};
}
// Standard Gumbel noise from 32 random bits: g = -log(E), E = -log(1-w) ~ Exp(1),
// w = (bits+0.5)*2^-32.  log1p keeps the E -> 0 tail (the large-g tail that decides rare
// wins) at full 32-bit resolution: g in [-2.8, 22.9].
__device__ __forceinline__ float gumbel_from_bits(uint32_t bits) {
    float w = (static_cast<float>(bits >> 1) + 0.5f) * 4.656612873077393e-10f;  // 2^-31
    w = fminf(w, 0.99999994f);  // float(bits>>1) can round up to 2^31
    const float e = -log1pf(-w);
    return -__logf(e);
}

// ---- reductions -----------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace b200vq

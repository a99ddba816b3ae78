// Shared device helpers for the b200vq kernels (sm_100a).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "b200vq.h"

namespace b200vq {

constexpr float kNormEps = 1e-12f;  // F.normalize eps (reference bottleneck.py:255,267)
constexpr int kMaxDim = 1024;

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
// ---- Gumbel noise, generated per GROUP of 8 consecutive codes --------------------------------
// The noise of code k for token row i is a pure function of (seed, offset, i, k) so draws do
// not depend on tiling or sharding.  It is built hierarchically so that a whole group can be
// ruled out from 16 random bits (the tensor-core sampler does that, vq_tc_kernel.cuh):
//   hi16    = halfword (group & 7) of Philox(stream 0; group >> 3, row)
//   a, b    = Philox(stream 1 | 2; group, row): words W0..W6 = a.xyzw, b.xyz ; b.w -> J, lo16
//   E_M     = Exp(1) variate from the 32 bits (hi16 << 16 | lo16);  E_M >= hi16 / 2^16
//   M       = -log(E_M / 8)      the largest of the group's 8 iid Gumbels ~ Gumbel(log 8)
//   J       = b.w >> 29          its position (uniform)
//   g_p     = -log(E_M / 8 + E_p), E_p ~ Exp(1) from the next unused word, p != J:
//             the other seven, iid Gumbel conditioned on being < M
// (max of n iid Gumbels is Gumbel(log n) at a uniform position, the rest are Gumbels
// truncated below it; P(g <= x | g <= M) = exp(-(e^-x - e^-M)).)
constexpr uint32_t kNoiseGroup = 8;
constexpr float kLn2 = 0.6931471805599453f;

// Exp(1) variate from 32 random bits, full relative resolution near 0 (the large-Gumbel tail).
// Always >= w = (bits + 0.5) / 2^32 >= (bits >> 16) / 2^16.
__device__ __forceinline__ float exp1_from_bits(uint32_t bits) {
    float w = fmaf(static_cast<float>(bits), 2.3283064365386963e-10f, 1.1641532182693481e-10f);  // (bits+.5)/2^32
    w = fminf(w, 0.99999994f);
    const float series = fmaf(w, 0.5f * w, w);             // -log(1-w) for tiny w
    const float direct = -kLn2 * __log2f(1.0f - w);
    return w < 9.765625e-4f ? series : direct;
}
__device__ __forceinline__ float neg_log(float e) { return -kLn2 * __log2f(e); }

__device__ __forceinline__ uint32_t noise_c3(uint64_t offset, uint32_t stream) {
    return (static_cast<uint32_t>(offset >> 32) & 0x3FFFFFFFu) | (stream << 30);
}
// the sixteen hi16 halfwords of groups 8*oct .. 8*oct+7 (two per word, low half first)
__device__ __forceinline__ Philox4 group_head_words(uint32_t group_oct, uint32_t grow, uint64_t seed, uint64_t offset) {
    return philox4x32_10(group_oct, grow, static_cast<uint32_t>(offset), noise_c3(offset, 0),
                         static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
}
__device__ __forceinline__ uint32_t pick_word(const Philox4& r, uint32_t i) {
    return i == 0 ? r.x : i == 1 ? r.y : i == 2 ? r.z : r.w;
}
__device__ __forceinline__ uint32_t head_hi16(const Philox4& r, uint32_t group) {
    const uint32_t w = pick_word(r, (group & 7u) >> 1);
    return (group & 1u) ? (w >> 16) : (w & 0xFFFFu);
}
// all 8 Gumbels of one group (the exact kernels and the sampler's event path use this)
__device__ __forceinline__ void group_noise(uint32_t group, uint32_t grow, uint64_t seed, uint64_t offset,
                                            uint32_t hi16, float (&g)[8]) {
    const Philox4 a = philox4x32_10(group, grow, static_cast<uint32_t>(offset), noise_c3(offset, 1),
                                    static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    const Philox4 b = philox4x32_10(group, grow, static_cast<uint32_t>(offset), noise_c3(offset, 2),
                                    static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    const uint32_t J = b.w >> 29;
    const float em8 = 0.125f * exp1_from_bits((hi16 << 16) | (b.w & 0xFFFFu));
    const uint32_t w[8] = {a.x, a.x, a.y, a.z, a.w, b.x, b.y, b.z};  // w[p] = W[p-1] (p >= 1)
    const uint32_t v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.z};  // v[p] = W[p]   (p <= 6)
#pragma unroll
    for (int p = 0; p < 8; ++p) {
        const uint32_t word = static_cast<uint32_t>(p) < J ? v[p] : w[p];
        const float e = static_cast<uint32_t>(p) == J ? 0.f : exp1_from_bits(word);
        g[p] = neg_log(em8 + e);
    }
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

// b200vq -- one pass over grad_output for the backward of out_linear (reference models/bottleneck.py:
// 162-164 under autograd):
//     grad_a [n][d]  = sum_c g[n][c] * w[c][d]   (+ ga_add[n][d])        grad of 'regularized_z'
//     grad_w [c][d]  = sum_n g[n][c] * a[n][d]                            out_linear.weight.grad
//     grad_b [c]     = sum_n g[n][c]                                      out_linear.bias.grad
// g (N x C, row stride ldg) is the only large operand (4*C bytes per token); linear_in_kernel +
// weight_grad_kernel read it twice, this kernel reads it once.  A CTA (16 warps, one per SM,
// persistent) walks 32-token tiles; a tile is streamed as C/256 stages of 32 tokens x 256 columns
// through a 3-deep cp.async ring.  In every stage warp w
//   * owns columns [16w, 16w+16) for grad_w: one m16 tile (rows = columns of g), k = the 32 tokens,
//     accumulators for all C/256 stages stay in registers for the whole kernel;
//   * owns tokens [16 (w&1), +16) and columns [32 (w>>1), +32) for grad_a: k = its 32 columns,
//     accumulated over the tile's stages, then written per column group to shared memory and summed when the tile's rows are stored.
// The ring is XOR-swizzled (16-byte groups inside each 128-byte segment, keyed by the token's low bits)
// so that both fragment patterns -- (token = k, column = row) and (token = row, column = k) -- are free
// of bank conflicts without padding.  grad_w / grad_b leave the CTA as one slab of the workspace;
// weight_grad_reduce_kernel adds the slabs in a fixed order (no atomics, deterministic).
#pragma once
#include "vq_linear.cuh"

namespace b200vq {

constexpr int kGoThreads = 512;
constexpr int kGoTok = 32;        // tokens per tile
constexpr int kGoCols = 256;      // columns per stage (16 warps x 16)
constexpr int kGoRing = 3;        // stages in flight

template <int D>
__host__ __device__ constexpr size_t go_smem_bytes_wgrad() {   // WITH_GA = false: ring + a tiles only
    return (static_cast<size_t>(kGoRing) * kGoTok * kGoCols + kGoRing * kGoTok * wg_ystride<D>()) * 4;
}
template <int D>
__host__ __device__ constexpr size_t go_smem_bytes(int C) {
    // w [C][YS] | ring [3][32][256] (sized for fp32 g) | a tiles [3][32][YS] | grad_a partials [8 column groups][32][D]
    return (static_cast<size_t>(C) * wg_ystride<D>() + kGoRing * kGoTok * kGoCols + kGoRing * kGoTok * wg_ystride<D>() +
            8 * kGoTok * D) * 4;
}
// element of the streamed operand as fp32 (bf16 and fp16 values are exact tf32 operands: no rounding step)
__device__ __forceinline__ float go_f32(float v) { return v; }
__device__ __forceinline__ float go_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float go_f32(__half v) { return __half2float(v); }

__device__ __forceinline__ void cp_async16_zfill(void* smem_dst, const void* gmem_src, bool valid) {
    const int bytes = valid ? 16 : 0;     // src-size 0: the 16 bytes are zero-filled, nothing is read
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))),
                 "l"(gmem_src), "r"(bytes)
                 : "memory");
}

// swizzle of a ring row: XOR applied to the column index (elements) of local token row r.  It permutes whole
// 16-byte groups inside a 128-byte segment: 4-element groups for fp32, 8-element groups for the 2-byte types.
template <typename GT>
__device__ __forceinline__ int go_swz(int r) {
    return sizeof(GT) == 4 ? (((r & 3) << 3) | (r & 4)) : (((r & 3) << 4) | ((r & 4) << 1));
}

// WITH_GA = false: weight / bias gradient only (in_linear.weight.grad = gz^T x with g = x, a = gz): no w, no ga.
template <int D, int NS, bool WITH_GA, typename GT = float>
__global__ void __launch_bounds__(kGoThreads, 1) grad_out_kernel(const GT* __restrict__ g, int64_t ldg, const float* __restrict__ a,
                                                                 const float* __restrict__ w, int64_t N,
                                                                 const float* __restrict__ ga_add, float* __restrict__ ga,
                                                                 float* __restrict__ partial) {
    constexpr int NT = D / 8, YS = wg_ystride<D>(), C = NS * kGoCols;
    extern __shared__ float s_go[];
    float* s_w = s_go;                                   // [C][YS] tf32 bits (WITH_GA only)
    GT* s_ring = reinterpret_cast<GT*>(s_w + (WITH_GA ? C * YS : 0));        // [3][32][256] swizzled
    float* s_a = s_w + (WITH_GA ? C * YS : 0) + kGoRing * kGoTok * kGoCols;  // [3][32][YS]
    constexpr int EPC = 16 / sizeof(GT);                 // elements per 16-byte cp.async
    float* s_ga = s_a + kGoRing * kGoTok * YS;           // [8][32][D]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, gq = lane >> 2, t = lane & 3;

    const int64_t tiles = (N + kGoTok - 1) / kGoTok;
    const int64_t mine = blockIdx.x < tiles ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;   // tiles of this CTA

    // stage s of this CTA's tile number j goes to ring slot `slot`; the tile's `a` rows ride with stage 0 (slot `aslot`).
    // All counters are 32-bit and advanced incrementally: 64-bit div / mod by NS or 3 per stage showed up in the SASS.
    auto issue = [&](int j, int s, int slot, int aslot) {
        const int64_t tok0 = (static_cast<int64_t>(blockIdx.x) + static_cast<int64_t>(j) * gridDim.x) * kGoTok;
        const int nrows = static_cast<int>(min(static_cast<int64_t>(kGoTok), N - tok0));   // >= 1
        const GT* src = g + tok0 * ldg + s * kGoCols;
        GT* dst = s_ring + slot * (kGoTok * kGoCols);
        constexpr int CPR = kGoCols / EPC;               // 16-byte groups per row
#pragma unroll
        for (int q = 0; q < kGoTok * CPR / kGoThreads; ++q) {
            const int i = threadIdx.x + q * kGoThreads, r = i / CPR, cg = i % CPR;
            const bool ok = r < nrows;
            cp_async16_zfill(dst + r * kGoCols + ((EPC * cg) ^ go_swz<GT>(r)), src + (ok ? r : 0) * ldg + EPC * cg, ok);
        }
        if (s == 0 && threadIdx.x < kGoTok * (D / 4)) {
            const int r = threadIdx.x / (D / 4), d4 = threadIdx.x - r * (D / 4);
            const bool ok = r < nrows;
            cp_async16_zfill(s_a + aslot * (kGoTok * YS) + r * YS + 4 * d4, a + (tok0 + (ok ? r : 0)) * D + 4 * d4, ok);
        }
    };

    float acc[NS][NT][4];     // grad_w: (column 256 s + 16 warp + gq [+8], d 8 j + 2 t [+1])
    float cs[NS][2];          // grad_b partial sums of columns 16 warp + gq, + 8 over this lane's tokens
    float gacc[NT][4];        // grad_a of the current tile over this warp's 32 columns of every stage
    constexpr bool AREG = NS * NT <= 9;   // room in the register file (C = 768, D = 24 yes; C = 1024, D = 32 no)
    uint32_t afrag[AREG ? kGoTok / 8 : 1][NT][2];   // tf32 B fragments of the current tile's `a` rows
#pragma unroll
    for (int s = 0; s < NS; ++s) {
        cs[s][0] = cs[s][1] = 0.f;
#pragma unroll
        for (int j = 0; j < NT; ++j) acc[s][j][0] = acc[s][j][1] = acc[s][j][2] = acc[s][j][3] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < NT; ++j) gacc[j][0] = gacc[j][1] = gacc[j][2] = gacc[j][3] = 0.f;

    // prologue: two stages in flight, then the weight matrix (overlaps the first loads)
    const int ntiles = static_cast<int>(mine);
    // (nj, ns, nslot, naslot): the next stage to issue
    int nj = 0, ns = 0, nslot = 0, naslot = 0;
    auto issue_next = [&]() {
        if (nj < ntiles) issue(nj, ns, nslot, naslot);
        cp_async_commit();
        nslot = nslot == kGoRing - 1 ? 0 : nslot + 1;
        if (++ns == NS) { ns = 0; ++nj; naslot = naslot == kGoRing - 1 ? 0 : naslot + 1; }
    };
    issue_next();
    issue_next();
    if constexpr (WITH_GA) {
        for (int i = threadIdx.x; i < C * D; i += kGoThreads) {
            const int c = i / D, d = i - c * D;
            s_w[c * YS + d] = __uint_as_float(tf32_rna(__ldg(w + i)));
        }
    }

    auto flush_ga = [&](int j) {   // tile j of this CTA: sum of the 8 column groups' partial rows -> global
        const int64_t tok0 = (static_cast<int64_t>(blockIdx.x) + static_cast<int64_t>(j) * gridDim.x) * kGoTok;
        const int nel = static_cast<int>(min(static_cast<int64_t>(kGoTok), N - tok0)) * D;
        for (int i = threadIdx.x; i < nel; i += kGoThreads) {
            float v = 0.f;
#pragma unroll
            for (int q = 0; q < 8; ++q) v += s_ga[q * (kGoTok * D) + i];
            const int64_t o = tok0 * D + i;
            ga[o] = v + (ga_add ? __ldg(ga_add + o) : 0.f);
        }
    };

    const int mt = warp & 1, kq = warp >> 1;
    const int swz_b = go_swz<GT>(gq);   // rows 16 mt + gq and + 8 share their low three bits
    const int sw0 = go_swz<GT>(t), sw1 = go_swz<GT>(t + 4);   // swizzles of token rows 8 ks + t and 8 ks + t + 4
    int slot = 0, aslot = 0;
    for (int j = 0; j < ntiles; ++j) {
        const float* ta = s_a + aslot * (kGoTok * YS);
#pragma unroll
        for (int s = 0; s < NS; ++s) {
            cp_async_wait<1>();
            __syncthreads();
            if (WITH_GA && s == 0 && j > 0) flush_ga(j - 1);
            issue_next();
            if (s == 0) {
                // the tile's `a` rows arrived with this stage: round them to tf32 once, in place, instead of in
                // every warp's fragment loads (all 16 warps read the whole tile in each of the NS stages)
                float* tw = s_a + aslot * (kGoTok * YS);
                for (int e = threadIdx.x; e < kGoTok * D; e += kGoThreads) {
                    const int r = e / D, d = e - r * D;
                    tw[r * YS + d] = __uint_as_float(tf32_rna(tw[r * YS + d]));
                }
                __syncthreads();
            }
            if (AREG && s == 0) {
                // B fragments of grad_w (the tile's `a` rows): the same for every stage of the tile, kept in registers
#pragma unroll
                for (int ks = 0; ks < kGoTok / 8; ++ks)
#pragma unroll
                    for (int n = 0; n < NT; ++n) {
                        afrag[ks][n][0] = __float_as_uint(ta[(8 * ks + t) * YS + 8 * n + gq]);
                        afrag[ks][n][1] = __float_as_uint(ta[(8 * ks + t + 4) * YS + 8 * n + gq]);
                    }
            }
            const GT* tile = s_ring + slot * (kGoTok * kGoCols);
            slot = slot == kGoRing - 1 ? 0 : slot + 1;
            // ---- grad_w / grad_b: rows = columns 16 warp + gq (+8), k = tokens
            const int col = 16 * warp + gq;
#pragma unroll
            for (int ks = 0; ks < kGoTok / 8; ++ks) {
                const GT* r0 = tile + (8 * ks + t) * kGoCols;             // token 8 ks + t
                const GT* r1 = r0 + 4 * kGoCols;                          // token 8 ks + t + 4
                const float x0 = go_f32(r0[col ^ sw0]), x1 = go_f32(r0[(col + 8) ^ sw0]);
                const float x2 = go_f32(r1[col ^ sw1]), x3 = go_f32(r1[(col + 8) ^ sw1]);
                cs[s][0] += x0 + x2;
                cs[s][1] += x1 + x3;
                const uint32_t a0 = tf32_op(x0), a1 = tf32_op(x1), a2 = tf32_op(x2), a3 = tf32_op(x3);
#pragma unroll
                for (int n = 0; n < NT; ++n) {
                    if constexpr (AREG) {
                        mma_tf32(acc[s][n], a0, a1, a2, a3, afrag[ks][n][0], afrag[ks][n][1]);
                    } else {
                        const uint32_t b0 = __float_as_uint(ta[(8 * ks + t) * YS + 8 * n + gq]);
                        const uint32_t b1 = __float_as_uint(ta[(8 * ks + t + 4) * YS + 8 * n + gq]);
                        mma_tf32(acc[s][n], a0, a1, a2, a3, b0, b1);
                    }
                }
            }
            if constexpr (!WITH_GA) continue;
            // ---- grad_a: rows = tokens 16 mt + gq (+8), k = columns 32 kq + 8 ks + t (+4) of this stage
            const GT* ra = tile + (16 * mt + gq) * kGoCols;
            const GT* rb = ra + 8 * kGoCols;
            const float* wb = s_w + (s * kGoCols + 32 * kq) * YS;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const int kc = 32 * kq + 8 * ks + t;   // (kc ^ swz_b) == ((32 kq + t) ^ swz_b) ^ (8 ks): bits 3-4 of 32 kq + t are 0
                const uint32_t a0 = tf32_op(go_f32(ra[kc ^ swz_b])), a1 = tf32_op(go_f32(rb[kc ^ swz_b]));
                const uint32_t a2 = tf32_op(go_f32(ra[(kc + 4) ^ swz_b])), a3 = tf32_op(go_f32(rb[(kc + 4) ^ swz_b]));
#pragma unroll
                for (int n = 0; n < NT; ++n) {
                    const uint32_t b0 = __float_as_uint(wb[(8 * ks + t) * YS + 8 * n + gq]);
                    const uint32_t b1 = __float_as_uint(wb[(8 * ks + t + 4) * YS + 8 * n + gq]);
                    mma_tf32(gacc[n], a0, a1, a2, a3, b0, b1);
                }
            }
            if (s == NS - 1) {
                if (NS == 1) __syncthreads();   // the flush of the previous tile (this iteration) reads the same buffer
                float* buf = s_ga + kq * (kGoTok * D) + (16 * mt + gq) * D + 2 * t;
#pragma unroll
                for (int n = 0; n < NT; ++n) {
                    *reinterpret_cast<float2*>(buf + 8 * n) = make_float2(gacc[n][0], gacc[n][1]);
                    *reinterpret_cast<float2*>(buf + 8 * D + 8 * n) = make_float2(gacc[n][2], gacc[n][3]);
                    gacc[n][0] = gacc[n][1] = gacc[n][2] = gacc[n][3] = 0.f;
                }
            }
        }
        aslot = aslot == kGoRing - 1 ? 0 : aslot + 1;
    }
    cp_async_wait<0>();
    __syncthreads();
    if (WITH_GA && ntiles > 0) flush_ga(ntiles - 1);

    // this CTA's slab: [C][D] then [C]
    float* slab = partial + static_cast<int64_t>(blockIdx.x) * (static_cast<int64_t>(C) * D + C);
#pragma unroll
    for (int s = 0; s < NS; ++s) {
        const int col = s * kGoCols + 16 * warp + gq;
#pragma unroll
        for (int n = 0; n < NT; ++n) {
            *reinterpret_cast<float2*>(slab + col * D + 8 * n + 2 * t) = make_float2(acc[s][n][0], acc[s][n][1]);
            *reinterpret_cast<float2*>(slab + (col + 8) * D + 8 * n + 2 * t) = make_float2(acc[s][n][2], acc[s][n][3]);
        }
        float c0 = cs[s][0], c1 = cs[s][1];
        c0 += __shfl_xor_sync(0xffffffffu, c0, 1); c0 += __shfl_xor_sync(0xffffffffu, c0, 2);
        c1 += __shfl_xor_sync(0xffffffffu, c1, 1); c1 += __shfl_xor_sync(0xffffffffu, c1, 2);
        if (t == 0) {
            slab[C * D + col] = c0;
            slab[C * D + col + 8] = c1;
        }
    }
}

}  // namespace b200vq

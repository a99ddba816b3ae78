// b200vq -- the Bottleneck's two skinny linears (in_linear C -> D, out_linear D -> C with
// D <= 32 << C; reference models/bottleneck.py:109-110, :140-142, :162-164) and their backward.
//
// These GEMMs move 4*C bytes per token for 2*C*D flops: they are HBM-bound by a wide margin
// (arithmetic intensity D/2 flop/byte), so they are written as streaming kernels that touch
// each activation byte exactly once, with the tiny weight matrix resident in shared memory.
// The contraction itself runs on warp-level tensor-core MMAs (mma.sync m16n8k8, TF32 inputs,
// fp32 accumulate) so that instruction issue never limits the stream:
//   * linear_in   (z = x W_in^T + b, decides the code indices) uses the 3xTF32 split
//     x = xb + xs, w = wb + ws, x.w ~ xb.wb + xb.ws + xs.wb  (error ~ 2^-21, fp32-class);
//   * linear_out / grad_x and the weight gradients use single-pass TF32 with round-to-nearest
//     operands (norm-wise error ~ 3e-4, inside the 1e-3 parity tolerance).
// Fragment layout of mma.sync.m16n8k8 (g = lane >> 2, t = lane & 3):
//   A: a0 (row g, k t)  a1 (row g+8, k t)  a2 (row g, k t+4)  a3 (row g+8, k t+4)
//   B: b0 (k t, col g)  b1 (k t+4, col g)
//   C: c0 (row g, col 2t)  c1 (row g, col 2t+1)  c2 (row g+8, col 2t)  c3 (row g+8, col 2t+1)
#pragma once
#include "vq_common.cuh"

namespace b200vq {

// fp32 -> tf32, round to nearest (ties away from zero).  sm_100a has no conversion instruction for this:
// cvt.rna.tf32.f32 is emulated as FSETP |v| >= inf ; @!p IADD 0x1000 ; LOP3 0xffffe000 (3 issue slots), and the
// streaming linears convert every operand element.
//   tf32_rna : the exact tf32 value, low 13 mantissa bits cleared (needed where it is subtracted from v)
//   tf32_op  : MMA operand only -- the tensor core ignores the low 13 bits, so the mask is dropped
//              (2 issue slots); NaN / Inf pass through unchanged like cvt.rna
__device__ __forceinline__ uint32_t tf32_rna(float v) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ uint32_t tf32_op(float v) {
    uint32_t b = __float_as_uint(v);
    if (fabsf(v) < __int_as_float(0x7f800000)) b += 0x1000u;   // false for NaN
    return b;
}
// 3xTF32 split: v = big + small exactly (fp32 subtraction of a tf32 from its source is exact); the small
// part goes to the tensor core truncated (that costs ~2^-21 of v, the size of the dropped small*small term)
__device__ __forceinline__ void tf32_split(float v, uint32_t& big, uint32_t& small) {
    big = tf32_rna(v);
    small = __float_as_uint(v - __uint_as_float(big));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// D (+)= A B with bf16 operands, k = 16.  Fragment layout (g = lane >> 2, t = lane & 3), two bf16 per register:
//   A: a0 (row g, k 2t..2t+1)  a1 (row g+8, same k)  a2 (row g, k 2t+8..2t+9)  a3 (row g+8, same k)
//   B: b0 (k 2t..2t+1, col g)  b1 (k 2t+8..2t+9, col g)          C: as m16n8k8
__device__ __forceinline__ void mma_bf16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))),
                 "l"(gmem_src)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

constexpr int kLinRows = 128;  // token rows per CTA (8 warps x 16)

// Output element types of the streaming linears: fp32, or the caller's autocast dtype (bf16 / fp16) when the module
// runs under torch.autocast -- the reference's nn.Linear hands back that dtype there (SURVEY.md 8b "dtype flow under AMP")
__device__ __forceinline__ void store2(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }
__device__ __forceinline__ void store2(__nv_bfloat16* p, float a, float b) { *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b); }
__device__ __forceinline__ void store2(__half* p, float a, float b) { *reinterpret_cast<__half2*>(p) = __floats2half2_rn(a, b); }
// four consecutive outputs, streaming (evict-first) store
__device__ __forceinline__ void store4_cs(float* p, float a, float b, float c, float d) { __stcs(reinterpret_cast<float4*>(p), make_float4(a, b, c, d)); }
__device__ __forceinline__ void store4_cs(__nv_bfloat16* p, float a, float b, float c, float d) {
    const __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
    __stcs(reinterpret_cast<uint2*>(p), make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi)));
}
__device__ __forceinline__ void store4_cs(__half* p, float a, float b, float c, float d) {
    const __half2 lo = __floats2half2_rn(a, b), hi = __floats2half2_rn(c, d);
    __stcs(reinterpret_cast<uint2*>(p), make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi)));
}

// =======================================================================================
// linear_in:  y[n][d] = sum_c x[n][c] * w[d][c] + b[d]      (x: N x C, row stride ldx; w: D x C, or C x D when w_is_cd)
// One warp = 16 rows.  x streams straight from global memory into A fragments (one float4
// per row per 16 columns: within a 16-column chunk lane t supplies columns 4t..4t+3, which
// the MMA sees as k = t, t+4 of two consecutive k8 steps; w is read with the same mapping).
// w lives in shared memory as fp32 (row stride C + 16 words: conflict-free LDS.128).
// =======================================================================================
constexpr int kLinInThreads = 512;                      // 16 warps share one staged weight matrix (1 CTA / SM)
constexpr int kLinInRows = kLinInThreads;               // linear_in: 32 rows per warp (two m16 tiles share every w fragment)

constexpr int kLinInRing = 4;   // RING: 16-column chunks in flight per row (thread-private cp.async slots in shared memory)
__host__ __device__ constexpr size_t lin_in_ring_bytes() { return static_cast<size_t>(kLinInRing) * 4 * kLinInThreads * 16; }

// RING = false: the next two chunks of a row wait in registers (8 x 16 bytes in flight per thread).  RING = true: a
// thread's next four chunks are in flight as cp.async into its own shared-memory slots (16 x 16 bytes per thread,
// 128 KB per SM) -- the kernel is bound by bytes in flight (one 512-row tile per CTA at 65 536 tokens: 24 dependent
// round trips per warp with two chunks in flight), not by the tensor pipe.  The slots are private to the thread that
// fills them, so no barrier is involved; used whenever the ring fits beside the weight matrix.
template <int D, bool PRECISE, typename OT = float, bool RING = false>
__global__ void __launch_bounds__(kLinInThreads, 1) linear_in_kernel(const float* __restrict__ x, int64_t ldx, int64_t N, int C,
                                                        const float* __restrict__ w, int w_is_cd,
                                                        const float* __restrict__ bias, OT* __restrict__ y, int ksplit = 1) {
    // ksplit (1, 2, 4, 8; RING only): small batches.  A CTA takes 512 / ksplit rows and its 16 warps split the columns
    // ksplit ways; the partial sums meet in shared memory and are added in a fixed order (deterministic).  A warp's time
    // is one pass over its columns whatever N is, so without the split 8 192 tokens keep 16 SMs busy for as long as
    // 65 536 keep 128.
    // the launch may use fewer than kLinInThreads threads (whole warps): a CTA tile is blockDim.x rows, so that the
    // number of tiles matches the number of SMs (65 536 tokens: 147 tiles of 448 rows instead of 128 of 512)
    const int nthr = static_cast<int>(blockDim.x);
    constexpr int NT = D / 8;
    extern __shared__ float s_w[];  // [D][C + 16]
    const int cs = C + 16;
    if (w_is_cd) {   // w given as [C][D] (out_linear.weight, for grad_a = grad_y W_out): transpose while staging
        for (int i = threadIdx.x; i < C * D; i += nthr) {
            const int c = i / D, d = i - c * D;
            s_w[d * cs + c] = __ldg(w + i);
        }
    } else {
        for (int i = threadIdx.x; i < D * (C / 4); i += nthr) {
            const int d = i / (C / 4), c4 = i - d * (C / 4);
            *reinterpret_cast<float4*>(&s_w[d * cs + 4 * c4]) = __ldg(reinterpret_cast<const float4*>(w + static_cast<int64_t>(d) * C) + c4);
        }
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int rpc = nthr / ksplit;                       // rows per CTA tile
    const int nrb = (nthr / 32) / ksplit;             // 32-row blocks per CTA tile
    const int rb = warp % nrb, sl = warp / nrb;                // this warp: row block, column slice
    for (int64_t tile = blockIdx.x; tile * rpc < N; tile += gridDim.x) {   // persistent: w is staged once per CTA
    const int64_t r0 = tile * rpc + rb * 32 + g;   // rows r0 + 8q, q = 0..3
    const float4* xr[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) xr[q] = reinterpret_cast<const float4*>(x + min(r0 + 8 * q, N - 1) * ldx) + t;
    float acc[2][NT][4];
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int j = 0; j < NT; ++j) { acc[m][j][0] = acc[m][j][1] = acc[m][j][2] = acc[m][j][3] = 0.f; }
    const int cbeg = sl * ((C / 16) / ksplit), chunks = cbeg + (C / 16) / ksplit;   // this warp's 16-column chunks [cbeg, chunks)
    constexpr int PF = RING ? 1 : 2;  // register path: chunks in flight per row
    float4 v[PF][4];
    float4* ring = reinterpret_cast<float4*>(s_w + D * cs) + threadIdx.x;   // slot (stage u, row q): ring[(4 u + q) * 512]
    if constexpr (RING) {
#pragma unroll
        for (int u = 0; u < kLinInRing - 1; ++u) {
            if (cbeg + u < chunks) {
#pragma unroll
                for (int q = 0; q < 4; ++q) cp_async16(ring + (4 * u + q) * nthr, xr[q] + 4 * (cbeg + u));
            }
            cp_async_commit();
        }
    } else {
#pragma unroll
        for (int u = 0; u < PF; ++u)
#pragma unroll
            for (int q = 0; q < 4; ++q) if (u < chunks) v[u][q] = __ldg(xr[q] + 4 * u);
    }
    for (int c0 = cbeg; c0 < chunks; c0 += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int c = c0 + u;
            if (c >= chunks) break;
            if constexpr (RING) {
                // refill the slot read in the previous iteration, then wait for this iteration's chunk
                const int nc = c + kLinInRing - 1;
                if (nc < chunks) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) cp_async16(ring + (4 * ((nc - cbeg) & (kLinInRing - 1)) + q) * nthr, xr[q] + 4 * nc);
                }
                cp_async_commit();
                cp_async_wait<kLinInRing - 1>();
#pragma unroll
                for (int q = 0; q < 4; ++q) v[0][q] = ring[(4 * ((c - cbeg) & (kLinInRing - 1)) + q) * nthr];
            }
            // PRECISE: x.w ~ xb.wb (TF32 x TF32, exact products) + xs.wb + xb.ws with the two correction products on
            // bf16 operands (k = 16 per MMA: 4 MMAs per 16 columns instead of 6).  xb is x as the tensor core sees it
            // (low 13 mantissa bits ignored), xs = x - xb exactly; the corrections are 2^-10 of the result, so their
            // bf16 operand rounding (2^-8) costs ~2^-19 of it: still fp32-class (referee tests: rel_err(z) < 2e-6).
            uint32_t xb[4][4];             // [row q][column 4t + e]: main operand bits
            uint32_t xsp[4][2], xbp[4][2]; // bf16 pairs (columns 4t, 4t+1 | 4t+2, 4t+3) of the small / big parts
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 a = v[u][q];
                if constexpr (!RING) { if (c + PF < chunks) v[u][q] = __ldg(xr[q] + 4 * (c + PF)); }
                if constexpr (PRECISE) {
                    const float e[4] = {a.x, a.y, a.z, a.w};
                    float sm[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        xb[q][i] = __float_as_uint(e[i]);
                        sm[i] = e[i] - __uint_as_float(xb[q][i] & 0xFFFFE000u);
                    }
                    xsp[q][0] = pack_bf16(sm[0], sm[1]); xsp[q][1] = pack_bf16(sm[2], sm[3]);
                    xbp[q][0] = pack_bf16(e[0], e[1]); xbp[q][1] = pack_bf16(e[2], e[3]);
                } else {
                    xb[q][0] = tf32_op(a.x); xb[q][1] = tf32_op(a.y); xb[q][2] = tf32_op(a.z); xb[q][3] = tf32_op(a.w);
                }
            }
#pragma unroll
            for (int j = 0; j < NT; ++j) {
                const float4 wv = *reinterpret_cast<const float4*>(&s_w[(8 * j + g) * cs + 16 * c + 4 * t]);
                if constexpr (PRECISE) {
                uint32_t wb[4], ws[4];
                tf32_split(wv.x, wb[0], ws[0]); tf32_split(wv.y, wb[1], ws[1]); tf32_split(wv.z, wb[2], ws[2]); tf32_split(wv.w, wb[3], ws[3]);
                const uint32_t wbp0 = pack_bf16(wv.x, wv.y), wbp1 = pack_bf16(wv.z, wv.w);
                const uint32_t wsp0 = pack_bf16(__uint_as_float(ws[0]), __uint_as_float(ws[1]));
                const uint32_t wsp1 = pack_bf16(__uint_as_float(ws[2]), __uint_as_float(ws[3]));
                // TF32 steps: step 0 = columns 4t, 4t+1 ; step 1 = columns 4t+2, 4t+3.  bf16 MMA: k 2t..2t+1 = columns
                // 4t, 4t+1 and k 2t+8..2t+9 = columns 4t+2, 4t+3 (the same column permutation on both operands).  The
                // tensor core adds into its accumulator with truncation, so every chunk starts from zero accumulators
                // (one for the big.big terms, one for the corrections) and the running sum is kept in fp32
                // round-to-nearest on the CUDA cores.
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    const int qa = 2 * m, qb = 2 * m + 1;
                    float cm[4] = {0.f, 0.f, 0.f, 0.f}, cc[4] = {0.f, 0.f, 0.f, 0.f};
                    mma_bf16(cc, xsp[qa][0], xsp[qb][0], xsp[qa][1], xsp[qb][1], wbp0, wbp1);
                    mma_bf16(cc, xbp[qa][0], xbp[qb][0], xbp[qa][1], xbp[qb][1], wsp0, wsp1);
                    mma_tf32(cm, xb[qa][0], xb[qb][0], xb[qa][1], xb[qb][1], wb[0], wb[1]);
                    mma_tf32(cm, xb[qa][2], xb[qb][2], xb[qa][3], xb[qb][3], wb[2], wb[3]);
#pragma unroll
                    for (int e = 0; e < 4; ++e) acc[m][j][e] += cm[e] + cc[e];
                }
                } else {   // single-pass TF32 (gradients: 1e-3 tolerance), accumulate in the tensor core
                    const uint32_t wb[4] = {tf32_op(wv.x), tf32_op(wv.y), tf32_op(wv.z), tf32_op(wv.w)};
#pragma unroll
                    for (int m = 0; m < 2; ++m) {
                        const int qa = 2 * m, qb = 2 * m + 1;
                        mma_tf32(acc[m][j], xb[qa][0], xb[qb][0], xb[qa][1], xb[qb][1], wb[0], wb[1]);
                        mma_tf32(acc[m][j], xb[qa][2], xb[qb][2], xb[qa][3], xb[qb][3], wb[2], wb[3]);
                    }
                }
            }
        }
    }
    if (!RING || ksplit == 1) {
#pragma unroll
    for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            const int d = 8 * j + 2 * t;
            const float b0 = bias ? __ldg(bias + d) : 0.f, b1 = bias ? __ldg(bias + d + 1) : 0.f;
            const int64_t ra = r0 + 16 * m, rbb = ra + 8;
            if (ra < N) store2(y + ra * D + d, acc[m][j][0] + b0, acc[m][j][1] + b1);
            if (rbb < N) store2(y + rbb * D + d, acc[m][j][2] + b0, acc[m][j][3] + b1);
        }
    if constexpr (RING) cp_async_wait<0>();
    } else {
        // column slices meet in shared memory (the ring's space: every thread's copies have landed and been read)
        cp_async_wait<0>();
        __syncthreads();
        float* part = s_w + D * cs;   // [ksplit][rpc][D]
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int j = 0; j < NT; ++j) {
                float* pr = part + (static_cast<size_t>(sl) * rpc + rb * 32 + g + 16 * m) * D + 8 * j + 2 * t;
                *reinterpret_cast<float2*>(pr) = make_float2(acc[m][j][0], acc[m][j][1]);
                *reinterpret_cast<float2*>(pr + 8 * D) = make_float2(acc[m][j][2], acc[m][j][3]);
            }
        __syncthreads();
        for (int i = threadIdx.x * 2; i < rpc * D; i += nthr * 2) {   // two consecutive outputs of one row (D is even)
            const int r = i / D, d = i - r * D;
            float s0 = 0.f, s1 = 0.f;
            for (int k = 0; k < ksplit; ++k) {
                const float2 pv = *reinterpret_cast<const float2*>(part + static_cast<size_t>(k) * rpc * D + i);
                s0 += pv.x; s1 += pv.y;
            }
            const int64_t row = tile * rpc + r;
            if (row < N) store2(y + row * D + d, s0 + (bias ? __ldg(bias + d) : 0.f), s1 + (bias ? __ldg(bias + d + 1) : 0.f));
        }
        __syncthreads();   // the next tile's copies reuse the space
    }
    }   // tile
}

// =======================================================================================
// linear_out:  y[n][c] = sum_d a[n][d] * w'[c][d] + b[c]     (a: N x D; y: N x C, row stride ldy)
// w' is given either as [C][D] (w_is_dc = 0: out_linear.weight) or as [D][C] (w_is_dc = 1:
// in_linear.weight, for grad_x = grad_z W_in).  It is kept in shared memory as TF32,
// [C][D + 4] (conflict-free B fragments).  One warp = 16 rows x all C columns.
// =======================================================================================
constexpr int kLinOutThreads = 512;   // 16 warps share one staged weight matrix
template <int D, typename OT = float>
__global__ void __launch_bounds__(kLinOutThreads) linear_out_kernel(const float* __restrict__ a, int64_t N, const float* __restrict__ w,
                                                         int w_is_dc, const float* __restrict__ bias, int C,
                                                         OT* __restrict__ y, int64_t ldy, int dbg,
                                                         const int64_t* __restrict__ gidx = nullptr, int64_t gK = 0, int gl2 = 0,
                                                         const float* __restrict__ add = nullptr, int64_t add_rows = 0,
                                                         int* __restrict__ bad_flag = nullptr) {
    // Decode variant (gidx != nullptr): row r of the A operand is normalise(a[gidx[r]]) -- `a` is then the
    // (gK, D) codebook -- and add[(r % add_rows)][:] (a positional table, nullable) joins the bias.
    constexpr int KS = D / 8, DS = D + 4;
    extern __shared__ float s_w[];  // [C][D + 4] tf32 bits, then bias [C]
    float* s_b = s_w + C * DS;
    for (int i = threadIdx.x; i < C; i += kLinOutThreads) s_b[i] = bias ? __ldg(bias + i) : 0.f;
    if (w_is_dc) {
        for (int i = threadIdx.x; i < C * D; i += kLinOutThreads) {
            const int d = i / C, c = i - d * C;
            s_w[c * DS + d] = __uint_as_float(tf32_rna(__ldg(w + i)));
        }
    } else {   // [C][D]: rows are contiguous, D % 4 == 0
        for (int i = threadIdx.x; i < C * (D / 4); i += kLinOutThreads) {
            const int c = i / (D / 4), q = i - c * (D / 4);
            const float4 v = __ldg(reinterpret_cast<const float4*>(w) + i);
            *reinterpret_cast<float4*>(&s_w[c * DS + 4 * q]) = make_float4(__uint_as_float(tf32_rna(v.x)), __uint_as_float(tf32_rna(v.y)),
                                                                            __uint_as_float(tf32_rna(v.z)), __uint_as_float(tf32_rna(v.w)));
        }
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    // Column order inside a pair of n8 tiles is permuted so that a lane's two C fragments are four
    // CONSECUTIVE output columns (one 16-byte store per row): MMA column n of the first tile is
    // output column 4*(n/2) + n%2 of the 16-column group, of the second tile 4*(n/2) + 2 + n%2.
    const int wrow = 4 * (g >> 1) + (g & 1);
    // Persistent: the weight staging above is paid once per CTA.  Work unit = 16 rows x a quarter of
    // the columns, dealt round-robin to all warps of the grid (fine enough to balance the last round).
    const int cq = (C / 16 + 3) / 4;   // 16-column groups per unit
    const int64_t units = ((N + 15) / 16) * 4;
    for (int64_t u = static_cast<int64_t>(blockIdx.x) * (kLinOutThreads / 32) + warp; u < units;
         u += static_cast<int64_t>(gridDim.x) * (kLinOutThreads / 32)) {
        const int64_t r0 = (u >> 2) * 16 + g;
        const int j0 = static_cast<int>(u & 3) * cq, j1 = min(C / 16, j0 + cq);
        int64_t ra = min(r0, N - 1), rb = min(r0 + 8, N - 1);
        uint32_t fa[KS][4];
        if (gidx) {
            ra = __ldg(gidx + ra); rb = __ldg(gidx + rb);
            const bool bada = ra < 0 || ra >= gK, badb = rb < 0 || rb >= gK;
            if (bada) ra = 0;
            if (badb) rb = 0;
            float va[2 * KS], vb[2 * KS], na = 0.f, nb = 0.f;
#pragma unroll
            for (int s = 0; s < KS; ++s) {
                va[2 * s] = __ldg(a + ra * D + 8 * s + t); va[2 * s + 1] = __ldg(a + ra * D + 8 * s + t + 4);
                vb[2 * s] = __ldg(a + rb * D + 8 * s + t); vb[2 * s + 1] = __ldg(a + rb * D + 8 * s + t + 4);
                na = fmaf(va[2 * s], va[2 * s], fmaf(va[2 * s + 1], va[2 * s + 1], na));
                nb = fmaf(vb[2 * s], vb[2 * s], fmaf(vb[2 * s + 1], vb[2 * s + 1], nb));
            }
            na += __shfl_xor_sync(0xffffffffu, na, 1); na += __shfl_xor_sync(0xffffffffu, na, 2);
            nb += __shfl_xor_sync(0xffffffffu, nb, 1); nb += __shfl_xor_sync(0xffffffffu, nb, 2);
            float sa = gl2 ? 1.f / fmaxf(sqrtf(na), 1e-12f) : 1.f, sb = gl2 ? 1.f / fmaxf(sqrtf(nb), 1e-12f) : 1.f;
            if (bada) { sa = __int_as_float(0x7fc00000); if (bad_flag && t == 0) *bad_flag = 1; }   // NaN row, as b200vq_decode
            if (badb) { sb = __int_as_float(0x7fc00000); if (bad_flag && t == 0) *bad_flag = 1; }
#pragma unroll
            for (int s = 0; s < KS; ++s) {
                fa[s][0] = tf32_rna(va[2 * s] * sa); fa[s][1] = tf32_rna(vb[2 * s] * sb);
                fa[s][2] = tf32_rna(va[2 * s + 1] * sa); fa[s][3] = tf32_rna(vb[2 * s + 1] * sb);
            }
        } else {
#pragma unroll
            for (int s = 0; s < KS; ++s) {
                fa[s][0] = tf32_rna(__ldg(a + ra * D + 8 * s + t));
                fa[s][1] = tf32_rna(__ldg(a + rb * D + 8 * s + t));
                fa[s][2] = tf32_rna(__ldg(a + ra * D + 8 * s + t + 4));
                fa[s][3] = tf32_rna(__ldg(a + rb * D + 8 * s + t + 4));
            }
        }
        const float* pa = nullptr;
        const float* pb = nullptr;
        if (add) {
            pa = add + (min(r0, N - 1) % add_rows) * C + 4 * t;
            pb = add + (min(r0 + 8, N - 1) % add_rows) * C + 4 * t;
        }
        OT* ya = y + r0 * ldy + 4 * t;
        OT* yb = y + (r0 + 8) * ldy + 4 * t;
        const bool oka = r0 < N, okb = r0 + 8 < N;
#pragma unroll 2
        for (int j = j0; j < j1; ++j) {
            float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
            const float* w0 = &s_w[(16 * j + wrow) * DS + t];
            const float* w1 = w0 + 2 * DS;
#pragma unroll
            for (int s = 0; s < KS; ++s) {
                if (dbg & 1) { acc0[0] += __uint_as_float(fa[s][0]); acc1[0] += __uint_as_float(fa[s][1]); continue; }
                mma_tf32(acc0, fa[s][0], fa[s][1], fa[s][2], fa[s][3], __float_as_uint(w0[8 * s]), __float_as_uint(w0[8 * s + 4]));
                mma_tf32(acc1, fa[s][0], fa[s][1], fa[s][2], fa[s][3], __float_as_uint(w1[8 * s]), __float_as_uint(w1[8 * s + 4]));
            }
            const float4 bb = *reinterpret_cast<const float4*>(&s_b[16 * j + 4 * t]);
            if (add) {
                const float4 ea = __ldg(reinterpret_cast<const float4*>(pa + 16 * j)), eb = __ldg(reinterpret_cast<const float4*>(pb + 16 * j));
                if (oka) store4_cs(ya + 16 * j, acc0[0] + bb.x + ea.x, acc0[1] + bb.y + ea.y, acc1[0] + bb.z + ea.z, acc1[1] + bb.w + ea.w);
                if (okb) store4_cs(yb + 16 * j, acc0[2] + bb.x + eb.x, acc0[3] + bb.y + eb.y, acc1[2] + bb.z + eb.z, acc1[3] + bb.w + eb.w);
                continue;
            }
            if (oka) store4_cs(ya + 16 * j, acc0[0] + bb.x, acc0[1] + bb.y, acc1[0] + bb.z, acc1[1] + bb.w);
            if (okb) store4_cs(yb + 16 * j, acc0[2] + bb.x, acc0[3] + bb.y, acc1[2] + bb.z, acc1[3] + bb.w);
        }
    }
}

// =======================================================================================
// linear_out, column-stationary: every warp of the CTA owns G consecutive 16-column groups for the whole kernel and keeps
// their B fragments (TF32) and bias IN REGISTERS; all 16 warps walk the same 16-row tiles.  Per row tile and group a warp
// issues 2 * D/8 MMAs and two 16-byte stores and nothing else -- the shared-memory kernel above spends 12 LDS on the same
// 6 MMAs, and LSU issue is what bounds it (a bf16 output, half the bytes, takes as long as fp32).  No shared memory, no
// weight staging: a small batch does not pay for it either.  C <= 16 * 16 * G columns, G <= 4.
// Contraction index permutation: MMA k = t reads d = 8s + 2t and k = t + 4 reads d = 8s + 2t + 1 in BOTH operands, so a
// lane's A elements of a k8 step are one 8-byte load.
// =======================================================================================
template <int D, int G, typename OT>
__global__ void __launch_bounds__(kLinOutThreads, 1) linear_out_cols_kernel(const float* __restrict__ a, int64_t N,
                                                                            const float* __restrict__ w, int w_is_dc,
                                                                            const float* __restrict__ bias, int C,
                                                                            OT* __restrict__ y, int64_t ldy,
                                                                            const float* __restrict__ add, int64_t add_rows) {
    constexpr int KS = D / 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int ngroups = C / 16, j0 = warp * G;
    // output column (inside a 16-column group) behind MMA column g of the group's first n8 tile; the second tile: + 2
    // (a lane's two C fragments are then four CONSECUTIVE output columns: one 16-byte store per row)
    const int wrow = 4 * (g >> 1) + (g & 1);
    uint32_t wb[G][2][KS][2];
    float bs[G][4];
#pragma unroll
    for (int jj = 0; jj < G; ++jj) {
        const int j = j0 + jj;
        const bool valid = j < ngroups;
#pragma unroll
        for (int tile = 0; tile < 2; ++tile) {
            const int col = 16 * j + wrow + 2 * tile;
#pragma unroll
            for (int s = 0; s < KS; ++s) {
                const int d0 = 8 * s + 2 * t;
                float w0 = 0.f, w1 = 0.f;
                if (valid) {
                    w0 = w_is_dc ? __ldg(w + static_cast<int64_t>(d0) * C + col) : __ldg(w + static_cast<int64_t>(col) * D + d0);
                    w1 = w_is_dc ? __ldg(w + static_cast<int64_t>(d0 + 1) * C + col) : __ldg(w + static_cast<int64_t>(col) * D + d0 + 1);
                }
                wb[jj][tile][s][0] = tf32_rna(w0);
                wb[jj][tile][s][1] = tf32_rna(w1);
            }
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) bs[jj][e] = (valid && bias) ? __ldg(bias + 16 * j + 4 * t + e) : 0.f;
    }
    const int64_t tiles = (N + 15) / 16;
    for (int64_t rt = blockIdx.x; rt < tiles; rt += gridDim.x) {
        const int64_t r0 = rt * 16 + g;
        const int64_t ra = min(r0, N - 1), rb = min(r0 + 8, N - 1);
        uint32_t fa[KS][4];
#pragma unroll
        for (int s = 0; s < KS; ++s) {
            const float2 va = __ldg(reinterpret_cast<const float2*>(a + ra * D + 8 * s + 2 * t));
            const float2 vb = __ldg(reinterpret_cast<const float2*>(a + rb * D + 8 * s + 2 * t));
            fa[s][0] = tf32_rna(va.x); fa[s][1] = tf32_rna(vb.x); fa[s][2] = tf32_rna(va.y); fa[s][3] = tf32_rna(vb.y);
        }
        const bool oka = r0 < N, okb = r0 + 8 < N;
        const float* pa = add ? add + (ra % add_rows) * C + 4 * t : nullptr;
        const float* pb = add ? add + (rb % add_rows) * C + 4 * t : nullptr;
        OT* ya = y + r0 * ldy + 4 * t;
        OT* yb = y + (r0 + 8) * ldy + 4 * t;
#pragma unroll
        for (int jj = 0; jj < G; ++jj) {
            const int j = j0 + jj;
            if (j >= ngroups) break;
            float acc0[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int s = 0; s < KS; ++s) {
                mma_tf32(acc0, fa[s][0], fa[s][1], fa[s][2], fa[s][3], wb[jj][0][s][0], wb[jj][0][s][1]);
                mma_tf32(acc1, fa[s][0], fa[s][1], fa[s][2], fa[s][3], wb[jj][1][s][0], wb[jj][1][s][1]);
            }
            float oa[4] = {acc0[0] + bs[jj][0], acc0[1] + bs[jj][1], acc1[0] + bs[jj][2], acc1[1] + bs[jj][3]};
            float ob[4] = {acc0[2] + bs[jj][0], acc0[3] + bs[jj][1], acc1[2] + bs[jj][2], acc1[3] + bs[jj][3]};
            if (add) {
                const float4 ea = __ldg(reinterpret_cast<const float4*>(pa + 16 * j)), eb = __ldg(reinterpret_cast<const float4*>(pb + 16 * j));
                oa[0] += ea.x; oa[1] += ea.y; oa[2] += ea.z; oa[3] += ea.w;
                ob[0] += eb.x; ob[1] += eb.y; ob[2] += eb.z; ob[3] += eb.w;
            }
            if (oka) store4_cs(ya + 16 * j, oa[0], oa[1], oa[2], oa[3]);
            if (okb) store4_cs(yb + 16 * j, ob[0], ob[1], ob[2], ob[3]);
        }
    }
}

// =======================================================================================
// weight_grad:  dw[c][d] += sum_n x[n][c] * y[n][d]     colsum[c] += sum_n x[n][c]
// (x: N x C row stride ldx -- the activation or the upstream gradient; y: N x D.)
// Used for out_linear.weight.grad (x = grad_output, y = regularized_z; colsum = bias grad)
// and in_linear.weight.grad (x = the layer input, y = grad_z; written transposed, [d][c]).
// The output tile is [C][D] with tokens as the contraction: a CTA walks 128-column chunks
// (outer) and its share of 64-token tiles (inner); a warp owns 32 columns of the chunk and one
// half of the tile's tokens, so its 32 x D accumulators stay in registers across all of the CTA's
// token tiles and are flushed with one atomicAdd per element per chunk.  x tiles are staged through shared
// memory with cp.async (double buffered): each x byte is read from HBM once.
// =======================================================================================
constexpr int kWgCols = 128;            // columns per chunk (8 warps x 16)
constexpr int kWgTok = 64;              // tokens per tile
constexpr int kWgStride = kWgCols + 8;  // 136 = 8 mod 32: conflict-free transposed A fragments

template <int D>
__host__ __device__ constexpr int wg_ystride() { return (D % 32 == 8 || D % 32 == 24) ? D : D + 8; }
template <int D>
__host__ __device__ constexpr size_t wg_smem_bytes() {
    return (2 * static_cast<size_t>(kWgTok) * kWgStride + 2 * static_cast<size_t>(kWgTok) * wg_ystride<D>()) * 4;
}

template <int D>
__global__ void __launch_bounds__(256) weight_grad_kernel(const float* __restrict__ x, int64_t ldx, const float* __restrict__ y,
                                                          int64_t N, int C, float* __restrict__ dw, int dw_is_dc,
                                                          float* __restrict__ colsum, float* __restrict__ partial,
                                                          int dbg) {
    constexpr int NT = D / 8, YS = wg_ystride<D>();
    extern __shared__ float s_wg[];
    float* s_x = s_wg;                                  // [2][128 tok][136]
    float* s_y = s_wg + 2 * kWgTok * kWgStride;         // [2][128 tok][YS]   (tf32 bits after conversion)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int64_t tiles = (N + kWgTok - 1) / kWgTok;
    const int chunks = (C + kWgCols - 1) / kWgCols;

    auto stage = [&](int buf, int64_t tile, int chunk) {
        // x block: 128 tokens x 128 columns (zero beyond N / C), y tile: 128 tokens x D
        const int64_t tok0 = tile * kWgTok;
        const int col0 = chunk * kWgCols;
        float* dx = s_x + buf * kWgTok * kWgStride;
        for (int i = threadIdx.x; i < kWgTok * (kWgCols / 4); i += 256) {
            const int r = i / (kWgCols / 4), c4 = i - r * (kWgCols / 4);
            float* dst = dx + r * kWgStride + 4 * c4;
            if (tok0 + r < N && col0 + 4 * c4 < C) cp_async16(dst, x + (tok0 + r) * ldx + col0 + 4 * c4);
            else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        float* dy = s_y + buf * kWgTok * YS;
        for (int i = threadIdx.x; i < kWgTok * (D / 4); i += 256) {
            const int r = i / (D / 4), d4 = i - r * (D / 4);
            float* dst = dy + r * YS + 4 * d4;
            if (tok0 + r < N) cp_async16(dst, y + (tok0 + r) * D + 4 * d4);
            else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        cp_async_commit();
    };

    // Warp w: columns [32 * (w & 3), +32) of the chunk (two m16 tiles that share every y fragment) and token
    // half (w >> 2) of each 64-token tile.  Row r of an m16 tile is chunk column 2r (r < 8) / 2(r-8)+1, so
    // the fragment pairs (row g, row g+8) are adjacent words: one 8-byte shared-memory load each.
    const int cb = warp & 3, th = warp >> 2;
    for (int chunk = 0; chunk < chunks; ++chunk) {
        float acc[2][NT][4];
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int j = 0; j < NT; ++j) { acc[m][j][0] = acc[m][j][1] = acc[m][j][2] = acc[m][j][3] = 0.f; }
        float cs[2][2] = {{0.f, 0.f}, {0.f, 0.f}};  // column sums of columns 2g / 2g+1 of each m tile over this lane's tokens
        int buf = 0;
        int64_t tile = blockIdx.x;
        if (tile < tiles) stage(0, tile, chunk);
        for (; tile < tiles; tile += gridDim.x) {
            const int64_t next = tile + gridDim.x;
            if (next < tiles) { stage(buf ^ 1, next, chunk); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
            __syncthreads();
            const float* bx = s_x + buf * kWgTok * kWgStride + 32 * cb + 2 * g;
            const float* by = s_y + buf * kWgTok * YS;
#pragma unroll
            for (int ss = 0; ss < kWgTok / 16; ++ss) {
                const int s = th * (kWgTok / 16) + ss;
                uint32_t a[2][4];
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    const float2 lo = *reinterpret_cast<const float2*>(bx + (8 * s + t) * kWgStride + 16 * m);
                    const float2 hi = *reinterpret_cast<const float2*>(bx + (8 * s + t + 4) * kWgStride + 16 * m);
                    cs[m][0] += lo.x + hi.x;
                    cs[m][1] += lo.y + hi.y;
                    a[m][0] = tf32_op(lo.x); a[m][1] = tf32_op(lo.y); a[m][2] = tf32_op(hi.x); a[m][3] = tf32_op(hi.y);
                }
#pragma unroll
                for (int j = 0; j < NT; ++j) {
                    const uint32_t b0 = tf32_op(by[(8 * s + t) * YS + 8 * j + g]);
                    const uint32_t b1 = tf32_op(by[(8 * s + t + 4) * YS + 8 * j + g]);
                    mma_tf32(acc[0][j], a[0][0], a[0][1], a[0][2], a[0][3], b0, b1);
                    mma_tf32(acc[1][j], a[1][0], a[1][1], a[1][2], a[1][3], b0, b1);
                }
            }
            __syncthreads();
            buf ^= 1;
        }
        // flush this chunk: c0/c1 -> (column 2g, d 2t / 2t+1), c2/c3 -> (column 2g+1, ...)
        if (dbg & 1) { if (acc[0][0][0] == 123.456f) dw[0] = 1.f; continue; }
        if (partial) {
            // No atomics: the two token halves are combined through shared memory and the CTA's
            // [C][D] (+ [C] column sums) partial goes to its own slab of the workspace with plain stores;
            // weight_grad_reduce_kernel adds the slabs (deterministic summation order).
            float* sc = s_x;                               // the x ring is idle here: [128 cols][D] + [128]
            float* scs = sc + kWgCols * D;
            if (th == 1) {
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    const int cl = 32 * cb + 16 * m + 2 * g;
#pragma unroll
                    for (int j = 0; j < NT; ++j) {
                        *reinterpret_cast<float2*>(sc + cl * D + 8 * j + 2 * t) = make_float2(acc[m][j][0], acc[m][j][1]);
                        *reinterpret_cast<float2*>(sc + (cl + 1) * D + 8 * j + 2 * t) = make_float2(acc[m][j][2], acc[m][j][3]);
                    }
                    float c0 = cs[m][0], c1 = cs[m][1];
                    c0 += __shfl_xor_sync(0xffffffffu, c0, 1); c0 += __shfl_xor_sync(0xffffffffu, c0, 2);
                    c1 += __shfl_xor_sync(0xffffffffu, c1, 1); c1 += __shfl_xor_sync(0xffffffffu, c1, 2);
                    if (t == 0) { scs[cl] = c0; scs[cl + 1] = c1; }
                }
            }
            __syncthreads();
            if (th == 0) {
                float* slab = partial + static_cast<int64_t>(blockIdx.x) * (static_cast<int64_t>(C) * D + C);
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    const int cl = 32 * cb + 16 * m + 2 * g, ca = chunk * kWgCols + cl;
#pragma unroll
                    for (int j = 0; j < NT; ++j) {
                        const int d = 8 * j + 2 * t;
                        const float2 p0 = *reinterpret_cast<const float2*>(sc + cl * D + d);
                        const float2 p1 = *reinterpret_cast<const float2*>(sc + (cl + 1) * D + d);
                        if (ca < C) *reinterpret_cast<float2*>(slab + static_cast<int64_t>(ca) * D + d) = make_float2(acc[m][j][0] + p0.x, acc[m][j][1] + p0.y);
                        if (ca + 1 < C) *reinterpret_cast<float2*>(slab + static_cast<int64_t>(ca + 1) * D + d) = make_float2(acc[m][j][2] + p1.x, acc[m][j][3] + p1.y);
                    }
                    float c0 = cs[m][0], c1 = cs[m][1];
                    c0 += __shfl_xor_sync(0xffffffffu, c0, 1); c0 += __shfl_xor_sync(0xffffffffu, c0, 2);
                    c1 += __shfl_xor_sync(0xffffffffu, c1, 1); c1 += __shfl_xor_sync(0xffffffffu, c1, 2);
                    if (t == 0) {
                        float* scol = slab + static_cast<int64_t>(C) * D;
                        if (ca < C) scol[ca] = c0 + scs[cl];
                        if (ca + 1 < C) scol[ca + 1] = c1 + scs[cl + 1];
                    }
                }
            }
            __syncthreads();
            continue;
        }
#pragma unroll
        for (int m = 0; m < 2; ++m) {
            const int ca = chunk * kWgCols + 32 * cb + 16 * m + 2 * g, cc = ca + 1;
#pragma unroll
            for (int j = 0; j < NT; ++j) {
                const int d = 8 * j + 2 * t;
                if (ca < C) {
                    atomicAdd(dw + (dw_is_dc ? static_cast<int64_t>(d) * C + ca : static_cast<int64_t>(ca) * D + d), acc[m][j][0]);
                    atomicAdd(dw + (dw_is_dc ? static_cast<int64_t>(d + 1) * C + ca : static_cast<int64_t>(ca) * D + d + 1), acc[m][j][1]);
                }
                if (cc < C) {
                    atomicAdd(dw + (dw_is_dc ? static_cast<int64_t>(d) * C + cc : static_cast<int64_t>(cc) * D + d), acc[m][j][2]);
                    atomicAdd(dw + (dw_is_dc ? static_cast<int64_t>(d + 1) * C + cc : static_cast<int64_t>(cc) * D + d + 1), acc[m][j][3]);
                }
            }
            if (colsum) {
                float c0 = cs[m][0], c1 = cs[m][1];
                c0 += __shfl_xor_sync(0xffffffffu, c0, 1); c0 += __shfl_xor_sync(0xffffffffu, c0, 2);
                c1 += __shfl_xor_sync(0xffffffffu, c1, 1); c1 += __shfl_xor_sync(0xffffffffu, c1, 2);
                if (t == 0) {
                    if (ca < C) atomicAdd(colsum + ca, c0);
                    if (cc < C) atomicAdd(colsum + cc, c1);
                }
            }
        }
    }
}

// dw (and colsum) = sum over the CTA slabs written by weight_grad_kernel / grad_out_kernel; block = (64, 8): 8 slab ranges
constexpr int kWgRedY = 8;   // slab ranges per block
__global__ void __launch_bounds__(64 * kWgRedY) weight_grad_reduce_kernel(const float* __restrict__ partial, int nslabs, int C, int D,
                                                                 float* __restrict__ dw, int dw_is_dc, float* __restrict__ colsum) {
    __shared__ float s_part[kWgRedY][64];
    const int total = C * D + C;
    const int i = blockIdx.x * 64 + threadIdx.x;
    float acc = 0.f;
    if (i < total) {
#pragma unroll 8
        for (int b = threadIdx.y; b < nslabs; b += kWgRedY) acc += __ldg(partial + static_cast<int64_t>(b) * total + i);
    }
    s_part[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && i < total) {
        float v = 0.f;
#pragma unroll
        for (int q = 0; q < kWgRedY; ++q) v += s_part[q][threadIdx.x];
        if (i < C * D) {
            const int c = i / D, d = i - c * D;
            dw[dw_is_dc ? d * C + c : i] = v;
        } else if (colsum) {
            colsum[i - C * D] = v;
        }
    }
}

// column sums of a small N x D matrix (in_linear.bias.grad = sum_n grad_z[n][:])
__global__ void __launch_bounds__(256) colsum_small_kernel(const float* __restrict__ y, int64_t N, int D, float* __restrict__ out) {
    // thread (r, d): r = threadIdx.x / D strides over rows
    __shared__ float s_acc[32];
    if (threadIdx.x < 32) s_acc[threadIdx.x] = 0.f;
    __syncthreads();
    const int per = 256 / D, d = threadIdx.x % D, r = threadIdx.x / D;
    float acc = 0.f;
    if (r < per) {
        const int64_t step = static_cast<int64_t>(gridDim.x) * per;
        int64_t n = static_cast<int64_t>(blockIdx.x) * per + r;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;     // four loads in flight per thread
        for (; n + 3 * step < N; n += 4 * step) {
            a0 += __ldg(y + n * D + d); a1 += __ldg(y + (n + step) * D + d);
            a2 += __ldg(y + (n + 2 * step) * D + d); a3 += __ldg(y + (n + 3 * step) * D + d);
        }
        for (; n < N; n += step) a0 += __ldg(y + n * D + d);
        acc = (a0 + a1) + (a2 + a3);
        atomicAdd(&s_acc[d], acc);
    }
    __syncthreads();
    if (threadIdx.x < D) atomicAdd(out + threadIdx.x, s_acc[threadIdx.x]);
}

}  // namespace b200vq

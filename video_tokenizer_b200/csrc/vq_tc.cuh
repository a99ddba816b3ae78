// b200vq -- tcgen05 (5th-gen tensor core) scoring kernels for sm_100a.
//
// The latent x codebook^T contraction is compute-bound on CUDA cores at the shipped codebook
// sizes (profiles/r01a_summary.md: 2*K*D flop per token against ~12*D bytes), so the scores
// are produced by tcgen05.mma with the accumulator tile in TMEM, and never leave the SM:
//
//   * operands are SPLIT bf16: x = hi + lo with hi = bf16(x), lo = bf16(x - hi).  One MMA
//     with the contraction dimension tripled computes  z_hi.e_hi + z_hi.e_lo + z_lo.e_hi,
//     which differs from the fp32 dot product by at most 3*2^-18 * sum|z_d||e_d| <= 1.2e-5
//     for unit vectors (the dropped lo.lo term and the two 2^-18 split residuals);
//   * deterministic modes keep a running best and second-best approximate score per row.
//     A row whose gap is below kTcMargin (4e-5 > 2 * error bound) is re-scored by the exact
//     fp32 kernel (resolve_kernel), every other row's approximate argmax IS the exact argmax;
//     the result is therefore bit-identical to the exact-fp32 path (vq_scores.cuh);
//   * stochastic mode is a lazy Gumbel-max: one draw per group of 8 codes bounds the
//     group's best perturbed score (vq_common.cuh: group noise); a group is expanded (its 8
//     Gumbels generated) only if that bound beats the row's current best.
//
// Layout: one CTA = 128 token rows x a range of 256-code tiles.  Warps 0-15 are four epilogue
// groups; group g drains columns [128*(g&1), +128) of TMEM accumulator buffer g>>1 (2 x 256
// columns) for tiles i = (g>>1) mod 2.  Warp 16 streams codebook tiles global->shared with
// cp.async.bulk (TMA engine) through a ring of mbarrier-guarded stages, warps 17 and 18 issue
// the MMAs of even / odd tiles (the per-tile wait/commit latency of one issuing thread would
// otherwise exceed the MMA time of a tile).  The codebook is pre-arranged in global memory in exactly the shared-memory
// core-matrix order (prepare_codebook_tc_kernel), so a tile is ONE contiguous bulk copy and
// no tensor map is needed.
#pragma once
#include "vq_scores.cuh"

namespace b200vq {

constexpr int kTcRows = 128;       // token rows per CTA (UMMA M)
constexpr int kTcCodes = 256;      // codes per tile (UMMA N)
constexpr int kTcBufs = 2;         // TMEM accumulator buffers (2 x 256 columns)
constexpr int kTcStages = 3;       // codebook tiles in flight
constexpr int kTcEpiWarps = 16;    // 4 epilogue groups: group g drains columns [128*(g&1), +128) of buffer g>>1
constexpr int kTcThreads = 32 * (kTcEpiWarps + 3);  // + producer + 2 MMA issuers (one per accumulator buffer)
constexpr float kTcMargin = 4e-5f; // ambiguity window on approximate scores (see header)

// Split operands.  With P = D/8 16-byte chunks per part, the contraction runs over the virtual
// chunk sequence   A: zh[0..P) zh[0..P) zl[0..P) (0)     B: eh[0..P) el[0..P) eh[0..P) (x)
// two chunks (one K=16 MMA) at a time.  B is stored ONCE as eh[0..P) el[0..P): the third part
// re-reads eh through its own descriptor, and because 2P is even no MMA straddles the wrap.
__host__ __device__ constexpr int tc_parts(int D) { return D / 8; }
__host__ __device__ constexpr int tc_slabs(int D) { return (3 * tc_parts(D) + 1) / 2; }           // MMAs per tile
__host__ __device__ constexpr int tc_a_chunks(int D) { return 2 * tc_slabs(D); }
__host__ __device__ constexpr int tc_b_chunks(int D) { return 2 * tc_parts(D); }
__host__ __device__ constexpr size_t tc_tile_bytes(int D) { return static_cast<size_t>(kTcCodes) * tc_b_chunks(D) * 16; }
__host__ __device__ constexpr bool tc_dim_ok(int D) { return D == 8 || D == 16 || D == 24 || D == 32; }
// first B chunk of MMA j (the second is the next one)
__host__ __device__ constexpr int tc_b_chunk_of_slab(int D, int j) { return (2 * j) % (2 * tc_parts(D)); }

// ---- PTX wrappers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra WAIT_DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t"
        "}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, bf16 x bf16 -> fp32, issued by ONE thread
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on an mbarrier once every previously issued MMA of this thread has completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread l of the warp gets row (lane base + l)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
// 32 lanes x 64 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld64(uint32_t taddr, uint32_t (&r)[64]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
        "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
        "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]), "=r"(r[32]),
          "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]),
          "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]), "=r"(r[48]),
          "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]),
          "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, no-swizzle shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// core matrix = 8 rows x 16 bytes stored as 128 contiguous bytes; lbo = byte distance between
// the two 16-byte K-chunks of one K=16 MMA, sbo = byte distance between 8-row groups.
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    return static_cast<uint64_t>((smem_addr >> 4) & 0x3FFFu) | (static_cast<uint64_t>((lbo >> 4) & 0x3FFFu) << 16) |
           (static_cast<uint64_t>((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// instruction descriptor: c=f32 (bit 4), a=b=bf16 (bits 7, 10), both K-major, N>>3 at 17, M>>4 at 24
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}

// ---- split-bf16 helpers ----------------------------------------------------------------------
__device__ __forceinline__ void split_bf16(float v, uint16_t& hi, uint16_t& lo) {
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    const __nv_bfloat16 l = __float2bfloat16_rn(v - __bfloat162float(h));
    hi = __bfloat16_as_ushort(h);
    lo = __bfloat16_as_ushort(l);
}
// 16-byte chunk c of the A operand's virtual sequence zh[0..P) zh[0..P) zl[0..P) (0)
template <int D>
__device__ __forceinline__ uint4 a_chunk(const uint16_t (&hi)[D], const uint16_t (&lo)[D], int c) {
    constexpr int P = tc_parts(D);
    uint32_t w[4] = {0u, 0u, 0u, 0u};
    if (c < 3 * P) {
        const int part = c / P, q = c % P;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const uint32_t a = part < 2 ? hi[q * 8 + 2 * e] : lo[q * 8 + 2 * e];
            const uint32_t b = part < 2 ? hi[q * 8 + 2 * e + 1] : lo[q * 8 + 2 * e + 1];
            w[e] = a | (b << 16);
        }
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// =======================================================================================
// Codebook image for the tensor-core path: for tile t, chunk c in [0, 2P) (eh chunks then el
// chunks), code r of the tile, the 8 bf16 values live at ((t * 2P + c) * 256 + r) * 16 bytes
// -- the shared-memory core-matrix order, so a tile is one contiguous bulk copy.  One thread
// per code.  flags[0] is set when a code row is not unit-norm (a zero row after
// normalisation): argmin-distance then differs from argmax-cosine and the deterministic
// tensor-core pass re-scores every row exactly.
// =======================================================================================
template <int D>
__global__ void __launch_bounds__(kTcCodes) prepare_codebook_tc_kernel(const float* __restrict__ emb, int64_t K,
                                                                       uint4* __restrict__ image,
                                                                       unsigned int* __restrict__ flags) {
    constexpr int P = tc_parts(D);
    const int64_t code = static_cast<int64_t>(blockIdx.x) * kTcCodes + threadIdx.x;  // padded to whole tiles
    const int64_t tile = blockIdx.x;
    const int r = threadIdx.x;
    uint16_t hi[D], lo[D];
    float e2 = 0.f;
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float v = code < K ? emb[code * D + d] : 0.f;
        e2 = fmaf(v, v, e2);
        split_bf16(v, hi[d], lo[d]);
    }
    if (code < K && !(fabsf(e2 - 1.0f) <= 1e-5f)) atomicOr(flags, 1u);
#pragma unroll
    for (int c = 0; c < 2 * P; ++c) {
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int d = (c % P) * 8 + 2 * e;
            const uint32_t a = c < P ? hi[d] : lo[d];
            const uint32_t b = c < P ? hi[d + 1] : lo[d + 1];
            w[e] = a | (b << 16);
        }
        image[(tile * (2 * P) + c) * kTcCodes + r] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

}  // namespace b200vq

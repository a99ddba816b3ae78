// b200vq -- tcgen05 stochastic (SVQ) scoring kernel, second generation.
//
// Reference semantics: idx ~ multinomial(softmax(cos / T))  (models/bottleneck.py:275-280), drawn here as a
// Gumbel-max over the streamed scores with the per-(row, code) noise of vq_common.cuh, so the draw is
// bit-identical to the exact-fp32 kernel's (vq_scores.cuh) for the same (seed, offset).
//
// What changed against the first sampler (vq_tc_kernel.cuh, MODE_SAMPLE):
//   * the tensor core only FILTERS.  Scores come from ONE bf16 product (z_hi . e_hi: 2 MMAs per 256-code tile
//     at D = 24 instead of 5), whose error is at most 2^-8 for unit vectors (two 2^-9 roundings, Cauchy-Schwarz).
//     A group of 8 codes is skipped only if   (max score + 2^-8) / T + (largest Gumbel its 16 head bits allow)
//     cannot beat the row's best perturbed score so far; that test is one FFMA per group and one compare per 32
//     columns;
//   * a group that passes is only QUEUED: (group id, row, head bits) + its 8 approximate scores as fp16, 24 bytes,
//     slot from one shared-memory atomic, no ballots.  The warp expands 32 queued groups at a time, one per lane:
//     the bound test is repeated per CODE against the row's CURRENT best (stale entries die here for a few
//     instructions), and the codes that survive (usually one) are re-scored in exact fp32 from the fp32 codebook
//     (same FMA order as scores_kernel), perturbed and merged into the row's key.  The final value of every
//     candidate is therefore exact -- the approximate scores never reach the result;
//   * warm start.  A running maximum over K/8 groups sees ~e^(lambda*slack) * ln(K/8) near-records.  The first
//     `scout` tiles are first only SCANNED for the row's best approximate group (no tests); that one group is
//     expanded, which puts the row's threshold within a couple of nats of its final value; then the remaining
//     tiles are tested, and the scout tiles are run again (tested) at the end.  Costs scout/ntiles more MMA
//     work, removes most of the early candidates.
//
// CTA layout: 128 rows x a range of 256-code tiles.  A tile is scored as four 64-code quarters (UMMA N = 64) and TMEM
// holds EIGHT accumulator buffers of 64 columns: (tile parity) x (quarter).  Warps 0-15 are four epilogue groups: group
// g (warps 4g..4g+3; warp w = rows [32*(w&3), +32)) drains quarter g of every tile, alternating between its two
// buffers, so the quarter of the next tile is already accumulated when the group finishes the current one, and only
// the four warps of a group are coupled through a buffer.
// Warp 16 streams codebook tiles with cp.async.bulk through a ring of stages, warps 17 / 18 issue the MMAs of even /
// odd tiles; a stage is handed back to the producer by a tcgen05.commit (no epilogue warp involved).
#pragma once
#include "vq_tc_kernel.cuh"

namespace b200vq {

constexpr int kTsStages = 6;          // codebook tiles in flight
constexpr int kTsThreads = 32 * (kTcEpiWarps + 3);   // 16 epilogue warps + producer + 2 MMA issuers
constexpr int kTsBufs = 8;            // TMEM accumulator buffers (8 x 64 columns)
constexpr int kTsPart = 64;           // codes per quarter tile (UMMA N)
constexpr int kTsQueueSlots = 256;    // per epilogue warp (ring): <= 31 left over + <= 128 pushed per chunk
constexpr float kTsScoreEps = 4.2e-3f;    // 2^-8 + 2^-18 (bf16 operands, unit vectors) + 2^-12 (fp16 copy in the queue) + fp32 accumulation

template <int D> __host__ __device__ constexpr int ts_chunks() { return 2 * ((D + 15) / 16); }   // 16-byte K chunks incl. zero padding
template <int D> __host__ __device__ constexpr uint32_t ts_stage_bytes() { return ts_chunks<D>() * kTcCodes * 16; }
template <int D> __host__ __device__ constexpr uint32_t ts_copy_bytes() { return (D / 8) * kTcCodes * 16; }  // the image's e_hi chunks
template <int D> __host__ __device__ constexpr int ts_zstride() { return D + 4; }
template <int D>
__host__ __device__ constexpr size_t ts_smem_bytes() {
    return 1024 /*align slack*/ + static_cast<size_t>(kTcRows) * ts_chunks<D>() * 16 + kTsStages * ts_stage_bytes<D>() +
           static_cast<size_t>(kTcRows) * ts_zstride<D>() * 4 + static_cast<size_t>(kTcEpiWarps) * kTsQueueSlots * 24 +
           kTcRows * 8 + 512;
}

// log2 of a 16-bit integer (exact conversion through the 2^23 trick; 0 -> -inf)
__device__ __forceinline__ float lg2_u16(uint32_t h) { return lg2_approx(__uint_as_float(0x4B000000u | h) - 8388608.0f); }

struct TsParams {
    const void* z;
    int64_t z_row_stride;
    int64_t N;
    const uint8_t* image;     // split-bf16 codebook image (prepare_codebook_tc_kernel); only the e_hi chunks are read
    const float* emb;         // (K, D) fp32 normalised codebook rows of this slice (exact expansion)
    int64_t K;                // codes in the slice
    int64_t k_offset;         // global index of code 0 (multiple of 256)
    int tiles_per_split;
    int scout_tiles;          // tiles scanned for a warm-start group before testing (0 = none)
    // 1-D grid.  The first head_row_tiles row tiles are split `splits` ways over the codes (tiles_per_split each); the
    // remaining row tiles -- the last, partly filled wave of a large batch -- are split tail_splits ways
    // (tail_tiles_per_split each, tail_scout_tiles) so that the wave is short instead of mostly empty.
    int head_row_tiles, splits;
    int tail_splits, tail_tiles_per_split, tail_scout_tiles;
    int l2norm;
    float t_inv;
    uint64_t seed, offset;
    int64_t row_offset;
    unsigned long long* keys; // (N) merged with atomicMax
    unsigned int* n_events;   // nullable: += groups expanded (statistics)
    int debug;                // tuning aid: 1 = no MMA, 2 = no epilogue work, 64 = candidates dropped
    int flush_early, flush_late, early_iters;   // smallest batch drained at a tile end (first iterations / later)
    // Philox (seed, offset) read on the DEVICE when non-null (CUDA-graph replays then draw fresh noise): [0] seed,
    // [1] offset (the host code advances it with a one-thread kernel after this launch)
    const unsigned long long* philox_dev;
};

// ---- expand up to 32 queued groups, one per lane.  Per code of the group the bound test is repeated against the
// row's current best; for the codes that survive, the group's Gumbels are generated and the APPROXIMATE perturbed
// score is compared with the current best (most survivors of the bound die here: the largest Gumbel of a group
// rarely sits on its best code); what is left gets its exact fp32 score (the arithmetic of scores_kernel) and is
// merged into the row's key ----------------------------------------------------------------------------------
template <int D>
__device__ __noinline__ void ts_flush(const uint4* q_sc, const uint2* q_meta, uint32_t qhead, uint32_t n, uint32_t lane,
                                      uint32_t row_base, const float* s_z, unsigned long long* s_key,
                                      const float* __restrict__ emb, uint32_t k_offset, uint32_t k_end_global,
                                      uint32_t grow0, float t_inv, float tl2e, float slack19, uint64_t seed,
                                      uint64_t offset) {
    if (lane < n) {
        const uint32_t slot = (qhead + lane) & (kTsQueueSlots - 1);
        const uint2 meta = q_meta[slot];
        const uint4 sc = q_sc[slot];
        const uint32_t group = meta.x, rl = row_base + (meta.y & 31u), hi16 = meta.y >> 16;
        const float best_now = key_value(s_key[rl]);
        const float nthr = fmaf(best_now, kLog2e, -slack19);
        const float lh = lg2_u16(hi16);
        const uint32_t code0 = group * 8u;
        const uint32_t w[4] = {sc.x, sc.y, sc.z, sc.w};
        float sa[8];
        uint32_t live = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
            sa[2 * i] = f.x; sa[2 * i + 1] = f.y;
            live |= fmaf(f.x, tl2e, -lh) > nthr ? (1u << (2 * i)) : 0u;
            live |= fmaf(f.y, tl2e, -lh) > nthr ? (2u << (2 * i)) : 0u;
        }
        if (code0 + 8u > k_end_global) live &= (1u << (k_end_global - code0)) - 1u;   // codes beyond the slice
        if (live != 0u) {
            float g[8];
            group_noise(group, grow0 + rl, seed, offset, hi16, g);
            // approximate perturbed scores: only codes that can still reach the current best are worth exact scoring
            const float margin = kTsScoreEps * t_inv;
            uint32_t keep = 0;
#pragma unroll
            for (int t = 0; t < 8; ++t) keep |= (fmaf(sa[t], t_inv, g[t]) + margin >= best_now) ? (1u << t) : 0u;
            live &= keep;
            if (live != 0u) {
                float zn[D];
                const float* zr = s_z + rl * ts_zstride<D>();
#pragma unroll
                for (int d = 0; d < D; d += 4) {
                    const float4 t = *reinterpret_cast<const float4*>(zr + d);
                    zn[d] = t.x; zn[d + 1] = t.y; zn[d + 2] = t.z; zn[d + 3] = t.w;
                }
                const float* e = emb + static_cast<size_t>(code0 - k_offset) * D;
                float bv = -INFINITY;
                uint32_t bk = code0;
                while (live != 0u) {                     // ascending code order: ties keep the lowest index
                    const int pos = __ffs(live) - 1;
                    live &= live - 1u;
                    float acc = 0.f;
#pragma unroll
                    for (int d = 0; d < D; d += 4) {
                        const float4 ev = __ldg(reinterpret_cast<const float4*>(e + pos * D + d));
                        acc = fmaf(zn[d + 0], ev.x, acc);
                        acc = fmaf(zn[d + 1], ev.y, acc);
                        acc = fmaf(zn[d + 2], ev.z, acc);
                        acc = fmaf(zn[d + 3], ev.w, acc);
                    }
                    float gp = g[0];
#pragma unroll
                    for (int t = 1; t < 8; ++t) gp = pos == t ? g[t] : gp;
                    const float v = __fmaf_rn(acc, t_inv, gp);
                    if (v > bv) { bv = v; bk = code0 + pos; }
                }
                atomicMax(s_key + rl, make_key(bv, bk));
            }
        }
    }
    __syncwarp();
}

// mbarrier wait with a suspend-time hint: the thread sleeps in hardware until the phase completes or the hint
// expires, instead of spinning through try_wait / branch (which burns issue slots the epilogue warps need)
__device__ __forceinline__ void mbar_wait_hint(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "TS_WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@p bra TS_WAIT_DONE;\n\t"
        "bra TS_WAIT_LOOP;\n\t"
        "TS_WAIT_DONE:\n\t"
        "}" ::"r"(bar), "r"(parity), "r"(0x989680u) : "memory");
}

// =======================================================================================
// grid = (row tiles of 128, code splits); block = kTsThreads; dynamic smem = ts_smem_bytes<D>()
// =======================================================================================
// DBG = true compiles the tuning switches of TsParams::debug in (tools/ts_probe.py); the production instantiation has none
template <int D, typename T, bool DBG = false>
__global__ void __launch_bounds__(kTsThreads, 1) tc_sample_kernel(const TsParams p) {
    constexpr int CH = ts_chunks<D>();          // K chunks per operand row (incl. zero padding)
    constexpr int P = D / 8;                    // real chunks
    constexpr int SLABS = CH / 2;               // MMAs (K = 16) per quarter tile
    constexpr uint32_t A_BYTES = kTcRows * CH * 16;
    constexpr uint32_t B_BYTES = ts_stage_bytes<D>();
    constexpr uint32_t A_LBO = kTcRows * 16, B_LBO = kTcCodes * 16, SBO = 128;
    constexpr int ZS = ts_zstride<D>();

    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment by pointer arithmetic on the shared array (keeps the shared address space: LDS / STS / ATOMS)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sA = smem;
    uint8_t* sB = sA + A_BYTES;
    float* s_z = reinterpret_cast<float*>(sB + kTsStages * B_BYTES);
    uint4* s_qsc = reinterpret_cast<uint4*>(s_z + kTcRows * ZS);                      // queued groups: 8 fp16 scores
    uint2* s_qmeta = reinterpret_cast<uint2*>(s_qsc + kTcEpiWarps * kTsQueueSlots);     // (group, lane | head << 16)
    unsigned long long* s_key = reinterpret_cast<unsigned long long*>(s_qmeta + kTcEpiWarps * kTsQueueSlots);
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_key + kTcRows);  // full[S], empty[S], tfull[8], tempty[8]
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(s_bar + 2 * kTsStages + 2 * kTsBufs);
    uint32_t* s_qtail = s_tmem + 4;   // (kTcEpiWarps) queue tails

    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total_tiles = static_cast<int>((p.K + kTcCodes - 1) / kTcCodes);
    const int head_ctas = p.head_row_tiles * p.splits;
    const bool tail = static_cast<int>(blockIdx.x) >= head_ctas;
    const int tb = static_cast<int>(blockIdx.x) - head_ctas;
    const int row_tile = tail ? p.head_row_tiles + tb / p.tail_splits : static_cast<int>(blockIdx.x) % p.head_row_tiles;
    const int part = tail ? tb % p.tail_splits : static_cast<int>(blockIdx.x) / p.head_row_tiles;
    const int per = tail ? p.tail_tiles_per_split : p.tiles_per_split;
    const int64_t row0 = static_cast<int64_t>(row_tile) * kTcRows;
    const int tile_begin = part * per;
    const int ntiles = max(0, min(per, total_tiles - tile_begin));
    const int scout = min(tail ? p.tail_scout_tiles : p.scout_tiles, ntiles);
    const int iters = ntiles + scout;           // iteration j: tile j (scan only while j < scout), then tiles 0..scout-1 again
    const uint32_t first_code = static_cast<uint32_t>(p.k_offset) + tile_begin * kTcCodes;
    const uint64_t seed = p.philox_dev ? p.philox_dev[0] : p.seed, offset = p.philox_dev ? p.philox_dev[1] : p.offset;

    const uint32_t bar_full = smem_u32(s_bar), bar_empty = bar_full + 8 * kTsStages;
    const uint32_t bar_tfull = bar_empty + 8 * kTsStages, bar_tempty = bar_tfull + 8 * kTsBufs;

    // ---- prologue ---------------------------------------------------------------------------------
    if (threadIdx.x < kTcRows) {
        const int64_t row = row0 + threadIdx.x;
        float zn[D];
        float z2, zlen;
        if (row < p.N) {
            load_normalised_row<D, T>(static_cast<const T*>(p.z) + row * p.z_row_stride, p.l2norm, zn, z2, zlen);
        } else {
#pragma unroll
            for (int d = 0; d < D; ++d) zn[d] = 0.f;
        }
#pragma unroll
        for (int d = 0; d < D; d += 4)
            *reinterpret_cast<float4*>(s_z + threadIdx.x * ZS + d) = make_float4(zn[d], zn[d + 1], zn[d + 2], zn[d + 3]);
#pragma unroll
        for (int c = 0; c < CH; ++c) {
            uint32_t w[4] = {0u, 0u, 0u, 0u};
            if (c < P) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const uint32_t a = __bfloat16_as_ushort(__float2bfloat16_rn(zn[c * 8 + 2 * e]));
                    const uint32_t b = __bfloat16_as_ushort(__float2bfloat16_rn(zn[c * 8 + 2 * e + 1]));
                    w[e] = a | (b << 16);
                }
            }
            *reinterpret_cast<uint4*>(sA + c * A_LBO + threadIdx.x * 16) = make_uint4(w[0], w[1], w[2], w[3]);
        }
        s_key[threadIdx.x] = make_key(-INFINITY, first_code);
        if (threadIdx.x < kTcEpiWarps) s_qtail[threadIdx.x] = 0u;
    } else if (warp == kTcEpiWarps) {
        if (lane == 0) {
            for (int s = 0; s < kTsStages; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
            for (int b = 0; b < kTsBufs; ++b) { mbar_init(bar_tfull + 8 * b, 1); mbar_init(bar_tempty + 8 * b, 4); }
            fence_barrier_init();
        }
    } else if (warp == kTcEpiWarps + 1) {
        tmem_alloc(smem_u32(s_tmem), 512);
    }
    if constexpr (CH > P) {
        // zero K-padding chunks of every codebook stage (the bulk copies only write chunks [0, P))
        constexpr int PAD16 = (CH - P) * kTcCodes;   // 16-byte words per stage
        for (int i = threadIdx.x; i < kTsStages * PAD16; i += kTsThreads) {
            const int s = i / PAD16, o = i % PAD16;
            *reinterpret_cast<uint4*>(sB + s * B_BYTES + P * B_LBO + o * 16) = make_uint4(0u, 0u, 0u, 0u);
        }
    }
    fence_proxy_async();  // A tile / zero padding (generic-proxy stores) -> visible to the tensor core (async proxy)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;

    if (warp == kTcEpiWarps) {
        // ===== codebook tile producer: one bulk copy (the e_hi chunks of the tile) per iteration ==============
        if (lane == 0) {
            const uint8_t* src = p.image + static_cast<size_t>(tile_begin) * tc_tile_bytes(D);
            for (int j = 0; j < iters; ++j) {
                const int s = j % kTsStages;
                const int tile = j < ntiles ? j : j - ntiles;
                mbar_wait_hint(bar_empty + 8 * s, ((j / kTsStages) & 1) ^ 1);
                mbar_expect_tx(bar_full + 8 * s, ts_copy_bytes<D>());
                bulk_g2s(smem_u32(sB + s * B_BYTES), src + static_cast<size_t>(tile) * tc_tile_bytes(D), ts_copy_bytes<D>(),
                         bar_full + 8 * s);
            }
        }
    } else if (warp > kTcEpiWarps) {
        // ===== MMA issuers: warp 17 issues even iterations (buffers 0-3), warp 18 odd ones (buffers 4-7) ============
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(kTcRows, kTsPart);
            const uint32_t a0 = smem_u32(sA);
            const int par = warp - (kTcEpiWarps + 1);
            // (a polling issuer that lets every quarter advance on its own -- mbarrier.test_wait over the four quarters
            // instead of this fixed order -- measured the same 235 us: the quarter groups stay close enough together)
            for (int j = par; j < iters; j += 2) {
                const int s = j % kTsStages;
                mbar_wait_hint(bar_full + 8 * s, (j / kTsStages) & 1);
#pragma unroll
                for (int part = 0; part < 4; ++part) {
                    const int buf = par * 4 + part;
                    mbar_wait_hint(bar_tempty + 8 * buf, ((j >> 1) & 1) ^ 1);
                    tc_fence_after();
                    const uint32_t b0 = smem_u32(sB + s * B_BYTES) + part * (kTsPart * 16);
#pragma unroll
                    for (int k = 0; k < SLABS; ++k) {
                        if (DBG && (p.debug & 1)) break;
                        umma_bf16(tmem_base + buf * kTsPart, umma_desc(a0 + k * 2 * A_LBO, A_LBO, SBO),
                                  umma_desc(b0 + k * 2 * B_LBO, B_LBO, SBO), idesc, k > 0);
                    }
                    umma_commit(bar_tfull + 8 * buf);
                }
                umma_commit(bar_empty + 8 * s);   // all four quarters have read the stage once this fires
            }
        }
    } else {
        // ===== epilogue: group g = warps 4g..4g+3 drains quarter g of every tile; warp w = rows [32*(w&3), +32) =======
        const uint32_t grp = warp >> 2;
        const uint32_t row_base = 32u * (warp & 3);
        const uint32_t row_local = row_base + lane;
        uint4* q_sc = s_qsc + warp * kTsQueueSlots;
        uint2* q_meta = s_qmeta + warp * kTsQueueSlots;
        volatile uint32_t* qtail_p = s_qtail + warp;
        uint32_t qhead = 0;   // warp-uniform
        const uint32_t grow0 = static_cast<uint32_t>(p.row_offset + row0);
        const uint32_t grow = grow0 + row_local;
        const uint32_t k_off32 = static_cast<uint32_t>(p.k_offset);
        const uint32_t k_end_global = static_cast<uint32_t>(p.k_offset + p.K);
        const uint32_t t_lane = tmem_base + (row_base << 16) + grp * kTsPart;
        const float tl2e = p.t_inv * kLog2e;
        // candidate iff  m * tl2e - log2(head)  >  vstar * log2(e) - (19 + slack):   slack covers the score error
        // (kTsScoreEps / T in log2 units) and the lg2 / FFMA rounding of the test itself
        const float slack19 = 19.02f + kTsScoreEps * tl2e;
        auto read_nthr = [&]() { return fmaf(key_value(s_key[row_local]), kLog2e, -slack19); };
        auto flush = [&](uint32_t n) {
            ts_flush<D>(q_sc, q_meta, qhead, n, lane, row_base, s_z, s_key, p.emb, k_off32, k_end_global, grow0, p.t_inv, tl2e,
                        slack19, seed, offset);
            qhead += n;
        };
        float nthr = -INFINITY;      // "everything can matter" until the row has a key
        float sbest = -INFINITY;     // scan: best tagged score (position within its 32-column chunk in the low 5 bits)
        uint32_t scode = first_code; // ... and the first code of that chunk
        uint32_t va[32];
        int tested = 0;
        const uint32_t uiters = static_cast<uint32_t>(iters), uscout = static_cast<uint32_t>(scout);
        for (uint32_t j = 0; j < uiters; ++j) {
            const uint32_t buf = (j & 1u) * 4u + grp;
            const bool scan_only = j < uscout;
            const uint32_t tile = j < static_cast<uint32_t>(ntiles) ? j : j - ntiles;
            const uint32_t code0 = first_code + tile * kTcCodes + grp * kTsPart;          // global code id of column 0
            const uint32_t group0 = code0 >> 3;                                          // multiple of 8
            const int valid = static_cast<int>(min(static_cast<uint32_t>(kTsPart), k_end_global > code0 ? k_end_global - code0 : 0u));
            Philox4 r0{};
            if (!scan_only) {
                // before the accumulator is needed: 16 random bits for each of the 8 groups (one Philox call)
                r0 = group_head_words(group0 >> 3, grow, seed, offset);
                nthr = read_nthr();
            }
            mbar_wait_hint(bar_tfull + 8 * buf, (j >> 1) & 1);
            tc_fence_after();
            if (DBG && (p.debug & 2)) {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_tempty + 8 * buf);
                continue;
            }
#pragma unroll
            for (int ch = 0; ch < 2; ++ch) {
                tmem_ld32(t_lane + (j & 1u) * 256u + ch * 32, va);
                tmem_ld_wait32(va);
                if (ch == 1) {   // the warp's rows of the buffer are in registers: hand them back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_tempty + 8 * buf);
                }
                if ((ch + 1) * 32 > valid) {   // columns beyond the slice: finite filler while scanning (it is tagged), -inf when testing
                    const uint32_t fill = scan_only ? __float_as_uint(kTcMasked) : 0xFF800000u;
#pragma unroll
                    for (int c = 0; c < 32; ++c) if (ch * 32 + c >= valid) va[c] = fill;
                }
                if (scan_only) {
                    // best single code of the chunk: every score carries its column in its 5 low mantissa bits
#pragma unroll
                    for (int c = 0; c < 32; ++c) va[c] = (va[c] & ~31u) | static_cast<uint32_t>(c);
                    const float mc = fmaxf(fmaxf(max8(&va[0]), max8(&va[8])), fmaxf(max8(&va[16]), max8(&va[24])));
                    scode = mc > sbest ? code0 + 32u * ch : scode;
                    sbest = fmaxf(sbest, mc);
                } else {
                    const float m0 = max8(&va[0]), m1 = max8(&va[8]), m2 = max8(&va[16]), m3 = max8(&va[24]);
                    // log2 of the chunk's four heads: the test   head < 2^(m * tl2e + thr)   is one FFMA + compare per group
                    // (head 0 -> -inf: always a candidate)
                    const uint32_t w0 = pick_word(r0, 2 * ch), w1 = pick_word(r0, 2 * ch + 1);
                    const uint32_t h0 = w0 & 0xFFFFu, h1 = w0 >> 16, h2 = w1 & 0xFFFFu, h3 = w1 >> 16;
                    const float c0 = fmaf(m0, tl2e, -lg2_u16(h0)), c1 = fmaf(m1, tl2e, -lg2_u16(h1));
                    const float c2 = fmaf(m2, tl2e, -lg2_u16(h2)), c3 = fmaf(m3, tl2e, -lg2_u16(h3));
                    const bool ev = fmaxf(fmaxf(c0, c1), fmaxf(c2, c3)) > nthr;
                    if (__any_sync(0xffffffffu, ev) && !(DBG && (p.debug & 64))) {
                        if (ev) {
                            const uint32_t mk = (c0 > nthr ? 1u : 0u) | (c1 > nthr ? 2u : 0u) | (c2 > nthr ? 4u : 0u) |
                                                (c3 > nthr ? 8u : 0u);
                            uint32_t slot = atomicAdd(s_qtail + warp, __popc(mk));
                            const uint32_t hd[4] = {h0, h1, h2, h3};
#pragma unroll
                            for (int g = 0; g < 4; ++g) {
                                if (mk & (1u << g)) {
                                    const uint32_t sl = slot & (kTsQueueSlots - 1);
                                    uint32_t hw[4];
#pragma unroll
                                    for (int i = 0; i < 4; ++i) {
                                        const __half2 h = __floats2half2_rn(__uint_as_float(va[8 * g + 2 * i]),
                                                                            __uint_as_float(va[8 * g + 2 * i + 1]));
                                        hw[i] = *reinterpret_cast<const uint32_t*>(&h);
                                    }
                                    q_sc[sl] = make_uint4(hw[0], hw[1], hw[2], hw[3]);
                                    q_meta[sl] = make_uint2(group0 + 4 * ch + g, lane | (hd[g] << 16));
                                    ++slot;
                                }
                            }
                        }
                        __syncwarp();
                        // Inside a quarter tile the ring is drained only when the next chunk could overflow it (<= 128
                        // pushes): while this warp expands groups its rows of the accumulator are not released.  No
                        // tcgen05.ld is in flight here, so the call may spill freely.
                        if (ch == 0 && *qtail_p - qhead > kTsQueueSlots - 128u) {
                            do { flush(32u); } while (*qtail_p - qhead > kTsQueueSlots - 160u);
                            nthr = read_nthr();
                        }
                    }
                }
            }
            if (scan_only) {
                if (j + 1 == uscout) {
                    // last scan: expand the row's best approximate code of this group's columns -> a
                    // near-final threshold for the row
                    const uint32_t code = scode + (__float_as_uint(sbest) & 31u), g = code >> 3, pos = code & 7u;
                    const Philox4 hw = group_head_words(g >> 3, grow, seed, offset);
                    const uint32_t slot = atomicAdd(s_qtail + warp, 1u) & (kTsQueueSlots - 1);
                    // scores: +inf for the chosen code (passes every test of the expansion), -inf for the other seven
                    uint32_t hw4[4] = {0xFC00FC00u, 0xFC00FC00u, 0xFC00FC00u, 0xFC00FC00u};
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (static_cast<int>(pos >> 1) == i) hw4[i] = (pos & 1u) ? 0x7C00FC00u : 0xFC007C00u;
                    q_sc[slot] = make_uint4(hw4[0], hw4[1], hw4[2], hw4[3]);
                    q_meta[slot] = make_uint2(g, lane | (head_hi16(hw, g) << 16));
                    __syncwarp();
                    flush(32u);
                }
            } else {
                // regular drain, off the accumulator's critical path
                const uint32_t fmin = static_cast<uint32_t>(tested < p.early_iters ? p.flush_early : p.flush_late);
                ++tested;
                __syncwarp();
                for (uint32_t pending = *qtail_p - qhead; pending != 0u && pending >= fmin; pending = *qtail_p - qhead)
                    flush(min(32u, pending));
            }
        }
        __syncwarp();
        for (uint32_t pending = *qtail_p - qhead; pending != 0u; pending = *qtail_p - qhead) flush(min(32u, pending));
        if (p.n_events && lane == 0) atomicAdd(p.n_events, qhead);
        // all four groups done -> publish the rows (named barrier over the 512 epilogue threads)
        asm volatile("bar.sync 1, 512;" ::: "memory");
        if (grp == 0) {
            const int64_t row = row0 + row_local;
            if (row < p.N && ntiles > 0) atomicMax(p.keys + row, s_key[row_local]);
        }
    }
    // ---- teardown ------------------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    if (warp == kTcEpiWarps + 1) tmem_dealloc(tmem_base, 512);
}

}  // namespace b200vq

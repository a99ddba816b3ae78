// b200vq -- the tcgen05 scoring kernel (see vq_tc.cuh for the design).
//
// Epilogue structure (both modes).  A thread owns one token row and sees its scores 32
// columns at a time (tcgen05.ld 32x32b.x32).  Per group of 8 codes it does a few branch-free
// instructions (3-input max tree + one compare) that decide whether the group can matter:
//   deterministic: group max within kTcMargin of the row's best so far;
//   stochastic:    16 random bits of the group's largest Gumbel say its bound can beat the
//                  row's best perturbed score so far.
// Groups that can matter are rare after the first few tiles, and rows of one warp hit them
// at different times, so they are not processed in place (that would idle 31 lanes): the
// lane pushes (group id, 8 scores) into its warp's shared-memory queue, and whenever 32
// entries are queued the warp processes them one per lane with all lanes busy, merging
// results into per-row keys in shared memory with atomics.
#pragma once
#include "vq_tc.cuh"

namespace b200vq {

// wait for outstanding tcgen05.ld and tie the destination registers to the wait so that the
// compiler cannot hoist their uses above it
__device__ __forceinline__ void tmem_ld_wait32(uint32_t (&r)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                   "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]),
                   "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]),
                   "+r"(r[30]), "+r"(r[31])
                 :
                 : "memory");
}

__device__ __forceinline__ float max8(const uint32_t* v) {
    const float a = fmaxf(fmaxf(__uint_as_float(v[0]), __uint_as_float(v[1])), __uint_as_float(v[2]));
    const float b = fmaxf(fmaxf(__uint_as_float(v[3]), __uint_as_float(v[4])), __uint_as_float(v[5]));
    const float c = fmaxf(__uint_as_float(v[6]), __uint_as_float(v[7]));
    return fmaxf(fmaxf(a, b), c);
}
__device__ __forceinline__ float ord_value(uint32_t o) {
    return __uint_as_float((o & 0x80000000u) ? (o ^ 0x80000000u) : ~o);
}
__device__ __forceinline__ float key_value(unsigned long long key) { return ord_value(static_cast<uint32_t>(key >> 32)); }
constexpr uint32_t kOrdNegInf = 0x007FFFFFu;                       // orderable(-inf)
constexpr unsigned long long kKeyNegInf = 0x007FFFFFull << 32;     // (-inf, no code)

constexpr int kQueueSlots = 64;   // per epilogue warp (ring); a push adds <= 32, a flush removes 32
constexpr int kQueueWords = 12;   // 8 scores, group id, row | hi16 << 16, 2 pad (48-byte entries)

struct TcParams {
    const void* z;
    int64_t z_row_stride;
    int64_t N;
    const uint4* image;       // split-bf16 codebook image (prepare_codebook_tc_kernel)
    int64_t K;                // codes in the image
    int64_t k_offset;         // global index of code 0 (multiple of 128)
    int tiles_per_split;
    int l2norm;
    float t_inv;
    uint64_t seed, offset;
    int64_t row_offset;
    unsigned long long* keys; // (N) merged with atomicMax
    float* second;            // (N) deterministic modes: runner-up approximate score (orderable bits, atomicMax)
    int debug;                // tuning aid (B200VQ_TC_DEBUG): 1 = no MMA, 2 = no epilogue work, 4 = no tile copies
};

// per-warp queue + per-row state shared by the epilogue warps of a CTA
struct Epi {
    float* queue;                 // this warp's ring
    unsigned long long* s_key;    // (128) per-row best (value, code)
    uint32_t* s_second;           // (128) per-row runner-up value, orderable bits (deterministic)
    uint32_t qhead, qtail;        // warp-uniform
    uint32_t row_local, lane;
    float thr;                    // lane's row: det = best - margin ; sample = 19.02 - vstar*log2(e)
    // stochastic only
    uint32_t grow0;               // global row id of local row 0
    float t_inv;
    uint64_t seed, offset;
    uint32_t k_end_global;
};

constexpr float kLog2e = 1.4426950408889634f;

// ---- deterministic: merge one queued group into its row's (best, runner-up) -----------------------
__device__ __noinline__ void det_flush(const float* queue, uint32_t qhead, uint32_t n, uint32_t lane,
                                       unsigned long long* s_key, uint32_t* s_second) {
    if (lane < n) {
        const float* e = queue + ((qhead + lane) & (kQueueSlots - 1)) * kQueueWords;
        const float4 s0 = *reinterpret_cast<const float4*>(e);
        const float4 s1 = *reinterpret_cast<const float4*>(e + 4);
        const uint2 meta = *reinterpret_cast<const uint2*>(e + 8);
        const float s[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
        float best = -INFINITY, sec = -INFINITY;
        uint32_t bi = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            sec = fmaxf(sec, fminf(s[j], best));
            if (s[j] > best) { best = s[j]; bi = j; }
        }
        const uint32_t rl = meta.y & 0xFFu;
        const unsigned long long old = atomicMax(s_key + rl, make_key(best, meta.x * 8 + bi));
        // the loser of every key comparison is a runner-up candidate
        const float runner = fmaxf(sec, fminf(key_value(old), best));
        atomicMax(s_second + rl, orderable(runner));
    }
    __syncwarp();
}

// ---- stochastic: expand one queued group (its 8 Gumbels) and merge its best perturbed score -------
__device__ __noinline__ void sample_flush(const float* queue, uint32_t qhead, uint32_t n, uint32_t lane,
                                          unsigned long long* s_key, uint32_t grow0, float t_inv, uint64_t seed,
                                          uint64_t offset, uint32_t k_end_global) {
    if (lane < n) {
        const float* e = queue + ((qhead + lane) & (kQueueSlots - 1)) * kQueueWords;
        const float4 s0 = *reinterpret_cast<const float4*>(e);
        const float4 s1 = *reinterpret_cast<const float4*>(e + 4);
        const uint2 meta = *reinterpret_cast<const uint2*>(e + 8);
        const uint32_t group = meta.x, rl = meta.y & 0xFFu, hi16 = meta.y >> 16;
        float g[8];
        group_noise(group, grow0 + rl, seed, offset, hi16, g);
        const float s[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
        float bv = -INFINITY;
        uint32_t bk = group * 8;
#pragma unroll
        for (int p = 0; p < 8; ++p) {
            const float v = fmaf(s[p], t_inv, g[p]);
            if (group * 8 + p < k_end_global && v > bv) { bv = v; bk = group * 8 + p; }
        }
        atomicMax(s_key + rl, make_key(bv, bk));
    }
    __syncwarp();
}

template <bool SAMPLE>
__device__ __forceinline__ void epi_flush(Epi& c, uint32_t n) {
    if (SAMPLE) {
        sample_flush(c.queue, c.qhead, n, c.lane, c.s_key, c.grow0, c.t_inv, c.seed, c.offset, c.k_end_global);
        c.thr = fmaf(key_value(c.s_key[c.row_local]), -kLog2e, 19.02f);
    } else {
        det_flush(c.queue, c.qhead, n, c.lane, c.s_key, c.s_second);
        c.thr = key_value(c.s_key[c.row_local]) - kTcMargin;
    }
    c.qhead += n;
}

// push the groups flagged in this warp (ballot b) and drain the queue when 32 are waiting
template <bool SAMPLE>
__device__ __forceinline__ void epi_push(Epi& c, uint32_t b, bool ev, const uint32_t* v8, uint32_t group, uint32_t hi16) {
    if (ev) {
        const uint32_t slot = (c.qtail + __popc(b & ((1u << c.lane) - 1u))) & (kQueueSlots - 1);
        float* e = c.queue + slot * kQueueWords;
        *reinterpret_cast<uint4*>(e) = make_uint4(v8[0], v8[1], v8[2], v8[3]);
        *reinterpret_cast<uint4*>(e + 4) = make_uint4(v8[4], v8[5], v8[6], v8[7]);
        *reinterpret_cast<uint2*>(e + 8) = make_uint2(group, c.row_local | (hi16 << 16));
    }
    c.qtail += __popc(b);
    __syncwarp();
    if (c.qtail - c.qhead >= 32u) epi_flush<SAMPLE>(c, 32u);
}

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// =======================================================================================
// grid = (row tiles of 128, code splits); block = 576 threads; dynamic smem = tc_smem_bytes<D>()
// =======================================================================================
template <int D>
__host__ __device__ constexpr size_t tc_smem_bytes() {
    return 1024 /*align slack*/ + static_cast<size_t>(kTcRows) * tc_a_chunks(D) * 16 + kTcStages * tc_tile_bytes(D) +
           kTcEpiWarps * kQueueSlots * kQueueWords * 4 + kTcRows * 8 + kTcRows * 4 + 256;
}

template <int D, typename T, int MODE>
__global__ void __launch_bounds__(kTcThreads, 1) tc_scores_kernel(const TcParams p) {
    constexpr int SLABS = tc_slabs(D);
    constexpr uint32_t A_BYTES = kTcRows * tc_a_chunks(D) * 16;
    constexpr uint32_t B_BYTES = static_cast<uint32_t>(tc_tile_bytes(D));
    constexpr uint32_t A_LBO = kTcRows * 16, B_LBO = kTcCodes * 16, SBO = 128;
    constexpr bool SAMPLE = MODE == B200VQ_MODE_SAMPLE;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sA = smem;
    uint8_t* sB = sA + A_BYTES;
    float* s_queue = reinterpret_cast<float*>(sB + kTcStages * B_BYTES);
    unsigned long long* s_key = reinterpret_cast<unsigned long long*>(s_queue + kTcEpiWarps * kQueueSlots * kQueueWords);
    uint32_t* s_second = reinterpret_cast<uint32_t*>(s_key + kTcRows);
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_second + kTcRows);  // full[S], empty[S], tfull[2], tempty[2]
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(s_bar + 2 * kTcStages + 2 * kTcBufs);

    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row0 = static_cast<int64_t>(blockIdx.x) * kTcRows;
    const int total_tiles = static_cast<int>((p.K + kTcCodes - 1) / kTcCodes);
    const int tile_begin = blockIdx.y * p.tiles_per_split;
    const int ntiles = min(p.tiles_per_split, total_tiles - tile_begin);
    const uint32_t first_code = static_cast<uint32_t>(p.k_offset) + tile_begin * kTcCodes;

    const uint32_t bar_full = smem_u32(s_bar), bar_empty = bar_full + 8 * kTcStages;
    const uint32_t bar_tfull = bar_empty + 8 * kTcStages, bar_tempty = bar_tfull + 8 * kTcBufs;

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
        uint16_t hi[D], lo[D];
#pragma unroll
        for (int d = 0; d < D; ++d) split_bf16(zn[d], hi[d], lo[d]);
#pragma unroll
        for (int c = 0; c < tc_a_chunks(D); ++c)
            *reinterpret_cast<uint4*>(sA + c * A_LBO + threadIdx.x * 16) = a_chunk<D>(hi, lo, c);
        s_key[threadIdx.x] = make_key(-INFINITY, first_code);
        s_second[threadIdx.x] = kOrdNegInf;
    } else if (warp == kTcEpiWarps) {
        if (lane == 0) {
            for (int s = 0; s < kTcStages; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
            for (int b = 0; b < kTcBufs; ++b) { mbar_init(bar_tfull + 8 * b, 1); mbar_init(bar_tempty + 8 * b, 256); }
            fence_barrier_init();
        }
    } else if (warp == kTcEpiWarps + 1) {
        tmem_alloc(smem_u32(s_tmem), 512);
    }
    fence_proxy_async();  // A tile (generic-proxy stores) -> visible to the tensor core (async proxy)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;

    if (warp == kTcEpiWarps) {
        // ===== codebook tile producer: one bulk copy per tile ========================================
        if (lane == 0) {
            const uint8_t* src = reinterpret_cast<const uint8_t*>(p.image) + static_cast<size_t>(tile_begin) * B_BYTES;
            for (int i = 0; i < ntiles; ++i) {
                const int s = i % kTcStages;
                mbar_wait(bar_empty + 8 * s, ((i / kTcStages) & 1) ^ 1);
                if (p.debug & 4) { mbar_arrive(bar_full + 8 * s); continue; }
                mbar_expect_tx(bar_full + 8 * s, B_BYTES);
                bulk_g2s(smem_u32(sB + s * B_BYTES), src + static_cast<size_t>(i) * B_BYTES, B_BYTES, bar_full + 8 * s);
            }
        }
    } else if (warp == kTcEpiWarps + 1) {
        // ===== MMA issuer =====================================================================================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(kTcRows, kTcCodes);
            for (int i = 0; i < ntiles; ++i) {
                const int buf = i % kTcBufs, s = i % kTcStages;
                mbar_wait(bar_tempty + 8 * buf, ((i / kTcBufs) & 1) ^ 1);
                mbar_wait(bar_full + 8 * s, (i / kTcStages) & 1);
                tc_fence_after();
                const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB + s * B_BYTES);
#pragma unroll
                for (int k = 0; k < SLABS; ++k) {
                    if (p.debug & 1) break;
                    umma_bf16(tmem_base + buf * kTcCodes, umma_desc(a0 + k * 2 * A_LBO, A_LBO, SBO),
                              umma_desc(b0 + tc_b_chunk_of_slab(D, k) * B_LBO, B_LBO, SBO), idesc, k > 0);
                }
                umma_commit(bar_empty + 8 * s);     // stage reusable once these MMAs have read it
                umma_commit(bar_tfull + 8 * buf);   // accumulator buffer ready for its epilogue groups
            }
        }
    } else {
        // ===== epilogue groups: group g = warps 4g..4g+3, columns [128*(g&1), +128) of buffer g>>1 ==========
        const int grp = warp >> 2, buf = grp >> 1, half = grp & 1;
        Epi c;
        c.queue = s_queue + warp * kQueueSlots * kQueueWords;
        c.s_key = s_key;
        c.s_second = s_second;
        c.qhead = c.qtail = 0;
        c.row_local = 32 * (warp & 3) + lane;
        c.lane = lane;
        c.thr = SAMPLE ? INFINITY : -INFINITY;   // "everything can matter" until the first flush
        c.grow0 = static_cast<uint32_t>(p.row_offset + row0);
        c.t_inv = p.t_inv;
        c.seed = p.seed;
        c.offset = p.offset;
        c.k_end_global = static_cast<uint32_t>(p.k_offset + p.K);
        const uint32_t t_lane = tmem_base + ((32u * (warp & 3)) << 16) + buf * kTcCodes + half * 128;
        const float tl2e = p.t_inv * kLog2e;
        uint32_t v[32];
        for (int i = buf; i < ntiles; i += kTcBufs) {
            const int64_t code0 = static_cast<int64_t>(tile_begin + i) * kTcCodes + half * 128;  // index within the image
            const uint32_t group0 = static_cast<uint32_t>((p.k_offset + code0) >> 3);            // multiple of 16
            const int valid = static_cast<int>(max(static_cast<int64_t>(0), min(static_cast<int64_t>(128), p.K - code0)));
            uint32_t hi16[16];
            if (SAMPLE) {   // 16 random bits per group, two Philox calls per tile, before the accumulator is needed
                const Philox4 r0 = group_head_words(group0 >> 3, c.grow0 + c.row_local, p.seed, p.offset);
                const Philox4 r1 = group_head_words((group0 >> 3) + 1, c.grow0 + c.row_local, p.seed, p.offset);
#pragma unroll
                for (int q = 0; q < 8; ++q) { hi16[q] = head_hi16(r0, q); hi16[8 + q] = head_hi16(r1, q); }
            }
            mbar_wait(bar_tfull + 8 * buf, (i / kTcBufs) & 1);
            tc_fence_after();
            if (p.debug & 2) { tc_fence_before(); mbar_arrive(bar_tempty + 8 * buf); continue; }
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
                tmem_ld32(t_lane + ch * 32, v);
                tmem_ld_wait32(v);
                if (ch == 3) {   // this group's half of the buffer is in registers: hand it back to the MMA warp
                    tc_fence_before();
                    mbar_arrive(bar_tempty + 8 * buf);
                }
                if ((ch + 1) * 32 > valid) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) if (ch * 32 + j >= valid) v[j] = 0xFF800000u;
                }
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const float m = max8(&v[8 * g]);
                    bool ev;
                    if (SAMPLE) {
                        const float fh = __uint_as_float(0x4B000000u | hi16[4 * ch + g]) - 8388608.0f;
                        ev = fh < ex2_approx(fmaf(m, tl2e, c.thr));
                    } else {
                        ev = m > c.thr;
                    }
                    const uint32_t b = __ballot_sync(0xffffffffu, ev);
                    if (b) epi_push<SAMPLE>(c, b, ev, &v[8 * g], group0 + 4 * ch + g, SAMPLE ? hi16[4 * ch + g] : 0u);
                }
            }
        }
        while (c.qtail != c.qhead) epi_flush<SAMPLE>(c, min(32u, c.qtail - c.qhead));
        // all four groups done -> publish the rows (named barrier over the 512 epilogue threads)
        asm volatile("bar.sync 1, 512;" ::: "memory");
        if (grp == 0) {
            const int64_t row = row0 + c.row_local;
            if (row < p.N && ntiles > 0) {
                const unsigned long long mine = s_key[c.row_local];
                const unsigned long long old = atomicMax(p.keys + row, mine);
                if (!SAMPLE) {
                    // cross-split merge: the loser of every key comparison is a runner-up candidate
                    float runner = ord_value(s_second[c.row_local]);
                    if (old != 0ull) runner = fmaxf(runner, fminf(key_value(old), key_value(mine)));
                    atomicMax(reinterpret_cast<unsigned int*>(p.second) + row, orderable(runner));
                }
            }
        }
    }
    // ---- teardown ------------------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    if (warp == kTcEpiWarps + 1) tmem_dealloc(tmem_base, 512);
}

// =======================================================================================
// Exact re-scoring of the rows the tensor-core pass could not decide (runner-up within
// kTcMargin of the best approximate score, or every row when the codebook has a non-unit
// row and the mode ranks by distance).  Same fp32 arithmetic as scores_kernel, so the final
// indices are bit-identical to the exact path.  Each CTA owns a contiguous range of rows,
// collects the undecided ones 256 at a time and re-scans the codebook for each with all
// 256 threads.
// =======================================================================================
template <int D, typename T, int MODE>
__global__ void __launch_bounds__(256) resolve_kernel(const T* __restrict__ z, int64_t z_row_stride, int64_t N,
                                                      const float* __restrict__ emb, const float* __restrict__ e2,
                                                      int64_t K, int64_t k_offset, int l2norm,
                                                      unsigned long long* __restrict__ keys,
                                                      const float* __restrict__ second,
                                                      const unsigned int* __restrict__ flags,
                                                      unsigned int* __restrict__ n_resolved) {
    __shared__ int s_list[256];
    __shared__ int s_count;
    __shared__ unsigned long long s_red[8];
    const bool all_rows = MODE == B200VQ_MODE_ARGMIN_DIST && flags != nullptr && (*flags & 1u);
    const int64_t per_cta = (N + gridDim.x - 1) / gridDim.x;
    const int64_t r_begin = static_cast<int64_t>(blockIdx.x) * per_cta;
    const int64_t r_end = min(N, r_begin + per_cta);
    for (int64_t base = r_begin; base < r_end; base += 256) {
        if (threadIdx.x == 0) s_count = 0;
        __syncthreads();
        const int64_t row = base + threadIdx.x;
        if (row < r_end) {
            const float best = key_value(keys[row]);
            const uint32_t so = reinterpret_cast<const unsigned int*>(second)[row];
            const float sec = so == 0u ? -INFINITY : __uint_as_float((so & 0x80000000u) ? (so ^ 0x80000000u) : ~so);
            if (all_rows || !(best - sec >= kTcMargin)) s_list[atomicAdd(&s_count, 1)] = threadIdx.x;
        }
        __syncthreads();
        const int cnt = s_count;
        for (int i = 0; i < cnt; ++i) {
            const int64_t r = base + s_list[i];
            float zn[D];
            float z2, zlen;
            load_normalised_row<D, T>(z + r * z_row_stride, l2norm, zn, z2, zlen);
            float best_v = -INFINITY;
            uint32_t best_k = static_cast<uint32_t>(k_offset);
#pragma unroll 4
            for (int64_t k = threadIdx.x; k < K; k += 256) {
                const float* e = emb + k * D;
                float acc = 0.f;
#pragma unroll
                for (int d = 0; d < D; d += 4) {
                    const float4 ev = __ldg(reinterpret_cast<const float4*>(e + d));
                    acc = fmaf(zn[d + 0], ev.x, acc);
                    acc = fmaf(zn[d + 1], ev.y, acc);
                    acc = fmaf(zn[d + 2], ev.z, acc);
                    acc = fmaf(zn[d + 3], ev.w, acc);
                }
                const float v = rank_value<MODE>(acc, z2, __ldg(e2 + k), 1.0f, 0.f);
                if (v > best_v) { best_v = v; best_k = static_cast<uint32_t>(k_offset + k); }
            }
            unsigned long long key = make_key(best_v, best_k);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
                key = other > key ? other : key;
            }
            if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = key;
            __syncthreads();
            if (threadIdx.x == 0) {
                unsigned long long m = s_red[0];
#pragma unroll
                for (int w = 1; w < 8; ++w) m = s_red[w] > m ? s_red[w] : m;
                keys[r] = m;
            }
            __syncthreads();
        }
        if (threadIdx.x == 0 && cnt && n_resolved) atomicAdd(n_resolved, static_cast<unsigned int>(cnt));
        __syncthreads();
    }
}

}  // namespace b200vq

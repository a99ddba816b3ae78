// b200vq -- the tcgen05 scoring kernel of the DETERMINISTIC modes (see vq_tc.cuh for the design; the stochastic
// mode has its own kernel, vq_tc_sample.cuh).
//
// Epilogue structure.  A thread owns one token row and sees its scores 32 columns at a time (tcgen05.ld 32x32b.x32).
// Per group of 8 codes it does a few branch-free instructions (3-input max tree, position tags in the low mantissa
// bits) that keep the row's best and second-best group maxima; the winning code is resolved afterwards in exact fp32.
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

__device__ __forceinline__ void tmem_ld_wait64(uint32_t (&r)[64]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                   "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]),
                   "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]),
                   "+r"(r[30]), "+r"(r[31]), "+r"(r[32]), "+r"(r[33]), "+r"(r[34]), "+r"(r[35]), "+r"(r[36]),
                   "+r"(r[37]), "+r"(r[38]), "+r"(r[39]), "+r"(r[40]), "+r"(r[41]), "+r"(r[42]), "+r"(r[43]),
                   "+r"(r[44]), "+r"(r[45]), "+r"(r[46]), "+r"(r[47]), "+r"(r[48]), "+r"(r[49]), "+r"(r[50]),
                   "+r"(r[51]), "+r"(r[52]), "+r"(r[53]), "+r"(r[54]), "+r"(r[55]), "+r"(r[56]), "+r"(r[57]),
                   "+r"(r[58]), "+r"(r[59]), "+r"(r[60]), "+r"(r[61]), "+r"(r[62]), "+r"(r[63])
                 :
                 : "memory");
}

__device__ __forceinline__ float max8(const uint32_t* v) {
    const float a = fmaxf(fmaxf(__uint_as_float(v[0]), __uint_as_float(v[1])), __uint_as_float(v[2]));
    const float b = fmaxf(fmaxf(__uint_as_float(v[3]), __uint_as_float(v[4])), __uint_as_float(v[5]));
    const float c = fmaxf(__uint_as_float(v[6]), __uint_as_float(v[7]));
    return fmaxf(fmaxf(a, b), c);
}
// Deterministic modes, branch-free.  Per chunk of 32 accumulator columns the lane forms the maxima of
// its four groups of 8 codes (3-input FMNMX trees), tags each with its position in the two low
// mantissa bits (a perturbation of <= 3 ulp, far inside kTcMargin) and takes the largest and
// second largest of the four.  The row keeps: the largest tagged group maximum seen, the first code
// of the chunk it came from, and the largest group maximum among all OTHER groups.  The winning
// code is found afterwards by exact fp32 scoring of the winning group's 8 codes
// (resolve_groups_kernel); rows whose two best groups are within kTcMargin are re-scored in full.
constexpr float kTcMasked = -1e30f;   // masked (invalid) columns: finite so that tagging cannot make a NaN
__device__ __forceinline__ float tag2(float v, uint32_t t) { return __uint_as_float((__float_as_uint(v) & ~3u) | t); }
struct DetTrack {
    float best, second;
    uint32_t chunk;
    __device__ __forceinline__ void add(const uint32_t* v, uint32_t chunk_code) {
        const float q0 = tag2(max8(v), 0u), q1 = tag2(max8(v + 8), 1u);
        const float q2 = tag2(max8(v + 16), 2u), q3 = tag2(max8(v + 24), 3u);
        const float a = fmaxf(q0, q1), b = fminf(q0, q1), c = fmaxf(q2, q3), d = fminf(q2, q3);
        const float m1 = fmaxf(a, c);
        const float m2 = fmaxf(fmaxf(fminf(a, c), b), d);
        second = fmaxf(fmaxf(second, fminf(m1, best)), m2);
        chunk = m1 > best ? chunk_code : chunk;
        best = fmaxf(best, m1);
    }
    __device__ __forceinline__ uint32_t code() const { return chunk + 8u * (__float_as_uint(best) & 3u); }
};
__device__ __forceinline__ void mask_tail32(uint32_t* v, int valid) {   // columns >= valid are no codes
#pragma unroll
    for (int j = 0; j < 32; ++j) if (j >= valid) v[j] = __float_as_uint(kTcMasked);
}
__device__ __forceinline__ float ord_value(uint32_t o) {
    return __uint_as_float((o & 0x80000000u) ? (o ^ 0x80000000u) : ~o);
}
__device__ __forceinline__ float key_value(unsigned long long key) { return ord_value(static_cast<uint32_t>(key >> 32)); }
constexpr uint32_t kOrdNegInf = 0x007FFFFFu;                       // orderable(-inf)
constexpr unsigned long long kKeyNegInf = 0x007FFFFFull << 32;     // (-inf, no code)

struct TcParams {
    const void* z;
    int64_t z_row_stride;
    int64_t N;
    const uint4* image;       // split-bf16 codebook image (prepare_codebook_tc_kernel)
    int64_t K;                // codes in the image
    int64_t k_offset;         // global index of code 0 (multiple of 256: whole tiles)
    int tiles_per_split;
    int l2norm;
    float t_inv;
    uint64_t seed, offset;
    int64_t row_offset;
    unsigned long long* keys; // (N) merged with atomicMax
    float* second;            // (N) deterministic modes: runner-up approximate score (orderable bits, atomicMax)
    int debug;                // tuning aid (B200VQ_TC_DEBUG): 1 = no MMA, 2 = no epilogue work, 4 = no tile copies
    unsigned int* n_events;   // unused (kept for layout compatibility of the tuning tools)
};

constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ float lg2_approx(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// =======================================================================================
// grid = (row tiles of 128, code splits); block = kTcThreads (608: 16 epilogue warps + producer + 2 MMA issuers); dynamic smem = tc_smem_bytes<D>()
// =======================================================================================
template <int D>
__host__ __device__ constexpr size_t tc_smem_bytes() {
    return 1024 /*align slack*/ + static_cast<size_t>(kTcRows) * tc_a_chunks(D) * 16 + kTcStages * tc_tile_bytes(D) +
           kTcRows * 8 + 2 * kTcRows * 4 + 512;
}

template <int D, typename T, int MODE>
__global__ void __launch_bounds__(kTcThreads, 1) tc_scores_kernel(const TcParams p) {
    constexpr int SLABS = tc_slabs(D);
    constexpr uint32_t A_BYTES = kTcRows * tc_a_chunks(D) * 16;
    constexpr uint32_t B_BYTES = static_cast<uint32_t>(tc_tile_bytes(D));
    constexpr uint32_t A_LBO = kTcRows * 16, B_LBO = kTcCodes * 16, SBO = 128;
    static_assert(MODE != B200VQ_MODE_SAMPLE, "stochastic mode: tc_sample_kernel (vq_tc_sample.cuh)");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sA = smem;
    uint8_t* sB = sA + A_BYTES;
    unsigned long long* s_key = reinterpret_cast<unsigned long long*>(sB + kTcStages * B_BYTES);
    uint32_t* s_second = reinterpret_cast<uint32_t*>(s_key + kTcRows);
    uint32_t* s_thr = s_second + kTcRows;
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_thr + kTcRows);  // full[S], empty[S], tfull[2], tempty[2]
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
        s_thr[threadIdx.x] = kOrdNegInf;
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
    } else if (warp >= kTcEpiWarps + 1) {
        // ===== MMA issuers: warp 17 issues the tiles of buffer 0, warp 18 those of buffer 1 ===============
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(kTcRows, kTcCodes);
            const int buf = warp - (kTcEpiWarps + 1);
            const uint32_t a0 = smem_u32(sA);
            for (int i = buf; i < ntiles; i += kTcBufs) {
                const int s = i % kTcStages;
                mbar_wait(bar_tempty + 8 * buf, ((i / kTcBufs) & 1) ^ 1);
                mbar_wait(bar_full + 8 * s, (i / kTcStages) & 1);
                tc_fence_after();
                const uint32_t b0 = smem_u32(sB + s * B_BYTES);
#pragma unroll
                for (int k = 0; k < SLABS; ++k) {
                    if (p.debug & 1) break;
                    umma_bf16(tmem_base + buf * kTcCodes, umma_desc(a0 + k * 2 * A_LBO, A_LBO, SBO),
                              umma_desc(b0 + tc_b_chunk_of_slab(D, k) * B_LBO, B_LBO, SBO), idesc, k > 0);
                }
                // one commit per tile: the epilogue group that sees it also frees the codebook stage
                umma_commit(bar_tfull + 8 * buf);
            }
        }
    } else {
        // ===== epilogue groups: group g = warps 4g..4g+3, columns [128*(g&1), +128) of buffer g>>1 ==========
        const int grp = warp >> 2, buf = grp >> 1, half = grp & 1;
        {
            // ---- deterministic modes: branch-free chunk tracking, two register buffers in flight ----------
            const uint32_t row_local = 32 * (warp & 3) + lane;
            const uint32_t t_lane = tmem_base + ((32u * (warp & 3)) << 16) + buf * kTcCodes + half * 128;
            DetTrack tr{kTcMasked, kTcMasked, first_code};
            uint32_t va[32], vb[32];
            for (int i = buf; i < ntiles; i += kTcBufs) {
                const int64_t code0 = static_cast<int64_t>(tile_begin + i) * kTcCodes + half * 128;
                const uint32_t chunk_code = static_cast<uint32_t>(p.k_offset + code0);
                const int valid = static_cast<int>(max(static_cast<int64_t>(0), min(static_cast<int64_t>(128), p.K - code0)));
                mbar_wait(bar_tfull + 8 * buf, (i / kTcBufs) & 1);
                tc_fence_after();
                if (half == 0 && (warp & 3) == 0 && lane == 0) mbar_arrive(bar_empty + 8 * (i % kTcStages));
                if (p.debug & 2) { tc_fence_before(); mbar_arrive(bar_tempty + 8 * buf); continue; }
                const bool ldon = !(p.debug & 16), mathon = !(p.debug & 8);
                if (ldon) tmem_ld32(t_lane, va);
                tmem_ld_wait32(va);
                if (ldon) tmem_ld32(t_lane + 32, vb);
                if (valid < 32) mask_tail32(va, valid);
                if (mathon) tr.add(va, chunk_code);
                tmem_ld_wait32(vb);
                if (ldon) tmem_ld32(t_lane + 64, va);
                if (valid < 64) mask_tail32(vb, valid - 32);
                if (mathon) tr.add(vb, chunk_code + 32);
                tmem_ld_wait32(va);
                if (ldon) tmem_ld32(t_lane + 96, vb);
                if (valid < 96) mask_tail32(va, valid - 64);
                if (mathon) tr.add(va, chunk_code + 64);
                tmem_ld_wait32(vb);
                tc_fence_before();
                mbar_arrive(bar_tempty + 8 * buf);   // this group's half of the buffer is in registers
                if (valid < 128) mask_tail32(vb, valid - 96);
                if (mathon) tr.add(vb, chunk_code + 96);
            }
            if (tr.best > 0.5f * kTcMasked) {
                const unsigned long long old = atomicMax(s_key + row_local, make_key(tr.best, tr.code()));
                // the loser of every key comparison is a runner-up candidate
                atomicMax(s_second + row_local, orderable(fmaxf(tr.second, fminf(key_value(old), tr.best))));
            }
            asm volatile("bar.sync 1, 512;" ::: "memory");
            if (grp == 0) {
                const int64_t row = row0 + row_local;
                if (row < p.N && ntiles > 0) {
                    const unsigned long long mine = s_key[row_local];
                    const unsigned long long old = atomicMax(p.keys + row, mine);
                    float runner = ord_value(s_second[row_local]);
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
// row and the mode ranks by distance).  resolve_groups_kernel lists their row ids,
// rescore_kernel re-scans the codebook for one listed row per CTA iteration with the same
// fp32 arithmetic as scores_kernel, so the final indices are bit-identical to the exact path.
// list[0] = count, list[4 ..] = row ids.
// =======================================================================================
// After tc_scores_kernel in a deterministic mode (64 rows per CTA; a thread per row decides and
// normalises, then a warp per row scores the chunk): keys[row] holds (best chunk
// maximum, first code of that chunk), second[row] the best maximum among the other chunks.
//   decided rows (gap >= kTcMargin): lane l scores code chunk+l in exact fp32 -- the same arithmetic as
//     scores_kernel / rescore_kernel -- and the warp's max-key (highest value, lowest index) replaces
//     keys[row];
//   undecided rows (or every row when the codebook has a non-unit row and the mode ranks by
//     distance): appended to list[] for rescore_kernel.
constexpr int kResolveRows = 64;   // rows per CTA of resolve_groups_kernel
template <int D, typename T, int MODE>
__global__ void __launch_bounds__(256) resolve_groups_kernel(const T* __restrict__ z, int64_t z_row_stride, int64_t N,
                                                             const float* __restrict__ emb, const float* __restrict__ e2,
                                                             int64_t K, int64_t k_offset, int l2norm,
                                                             unsigned long long* __restrict__ keys,
                                                             const float* __restrict__ second,
                                                             const unsigned int* __restrict__ flags,
                                                             unsigned int* __restrict__ list) {
    __shared__ __align__(16) float s_z[kResolveRows][D + 4];   // normalised row, then |z|^2
    __shared__ uint32_t s_code[kResolveRows];                   // first code of the winning chunk, or ~0: skip
    const int64_t row0 = static_cast<int64_t>(blockIdx.x) * kResolveRows;
    const bool all_rows = MODE == B200VQ_MODE_ARGMIN_DIST && flags != nullptr && (*flags & 1u);
    if (threadIdx.x < kResolveRows) {   // one thread per row: decide, normalise once
        const int64_t row = row0 + threadIdx.x;
        uint32_t code = 0xFFFFFFFFu;
        if (row < N) {
            const unsigned long long key = keys[row];
            const uint32_t so = reinterpret_cast<const unsigned int*>(second)[row];
            const float sec = so == 0u ? -INFINITY : ord_value(so);
            if (all_rows || !(key_value(key) - sec >= kTcMargin)) {
                list[4 + atomicAdd(list, 1u)] = static_cast<unsigned int>(row);
                keys[row] = 0ull;   // rescore_kernel merges its code segments with atomicMax
            } else {
                float zn[D];
                float z2, zlen;
                load_normalised_row<D, T>(z + row * z_row_stride, l2norm, zn, z2, zlen);
#pragma unroll
                for (int d = 0; d < D; ++d) s_z[threadIdx.x][d] = zn[d];
                s_z[threadIdx.x][D] = z2;
                code = key_code(key);
            }
        }
        s_code[threadIdx.x] = code;
    }
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int it = 0; it < kResolveRows / 32; ++it) {   // a warp takes 4 rows at a time, 8 lanes (codes) per row
        const int r = it * 32 + warp * 4 + (lane >> 3);
        const uint32_t code = s_code[r];
        const int64_t k = static_cast<int64_t>(code) - k_offset + (lane & 7);
        unsigned long long mine = 0ull;
        if (code != 0xFFFFFFFFu && k < K) {
            const float* e = emb + k * D;
            float acc = 0.f;
#pragma unroll
            for (int d = 0; d < D; d += 4) {
                const float4 ev = __ldg(reinterpret_cast<const float4*>(e + d));
                const float4 zv = *reinterpret_cast<const float4*>(&s_z[r][d]);
                acc = fmaf(zv.x, ev.x, acc);
                acc = fmaf(zv.y, ev.y, acc);
                acc = fmaf(zv.z, ev.z, acc);
                acc = fmaf(zv.w, ev.w, acc);
            }
            mine = make_key(rank_value<MODE>(acc, s_z[r][D], __ldg(e2 + k), 1.0f, 0.f), static_cast<uint32_t>(k_offset + k));
        }
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, mine, o);
            mine = other > mine ? other : mine;
        }
        if ((lane & 7) == 0 && code != 0xFFFFFFFFu) keys[row0 + r] = mine;
    }
}

constexpr int kRescoreSegs = 8;   // CTAs that share one listed row (each scans K / 8 codes)
template <int D, typename T, int MODE>
__global__ void __launch_bounds__(256) rescore_kernel(const T* __restrict__ z, int64_t z_row_stride,
                                                      const float* __restrict__ emb, const float* __restrict__ e2,
                                                      int64_t K, int64_t k_offset, int l2norm,
                                                      unsigned long long* __restrict__ keys,
                                                      const unsigned int* __restrict__ list,
                                                      unsigned int* __restrict__ n_resolved) {
    __shared__ unsigned long long s_red[8];
    const unsigned int count = list[0];
    if (blockIdx.x == 0 && threadIdx.x == 0 && n_resolved) atomicAdd(n_resolved, count);
    const unsigned int seg = blockIdx.x % kRescoreSegs;
    const int64_t per = (K + kRescoreSegs - 1) / kRescoreSegs;
    const int64_t k_begin = seg * per, k_end = min(K, k_begin + per);
    for (unsigned int it = blockIdx.x / kRescoreSegs; it < count; it += gridDim.x / kRescoreSegs) {
        const int64_t r = list[4 + it];
        float zn[D];
        float z2, zlen;
        load_normalised_row<D, T>(z + r * z_row_stride, l2norm, zn, z2, zlen);
        float best_v = -INFINITY;
        uint32_t best_k = static_cast<uint32_t>(k_offset + k_begin);
#pragma unroll 4
        for (int64_t k = k_begin + threadIdx.x; k < k_end; k += 256) {
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
        unsigned long long key = k_begin < k_end ? make_key(best_v, best_k) : 0ull;
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
            atomicMax(keys + r, m);   // resolve_groups_kernel zeroed the keys of listed rows
        }
        __syncthreads();
    }
}

}  // namespace b200vq

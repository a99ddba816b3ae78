// b200vq -- the tcgen05 scoring kernel (see vq_tc.cuh for the design).
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
__device__ __forceinline__ float key_value(unsigned long long key) {
    const uint32_t o = static_cast<uint32_t>(key >> 32);
    return __uint_as_float((o & 0x80000000u) ? (o ^ 0x80000000u) : ~o);
}
constexpr unsigned long long kKeyNegInf = 0x007FFFFFull << 32;  // orderable(-inf), no code

struct TcSmem {
    static constexpr int kQueueSlots = 64;   // per epilogue warp (ring)
    static constexpr int kQueueWords = 12;   // 8 scores, group, row, 2 pad
};

struct TcParams {
    const void* z;
    int64_t z_row_stride;
    int64_t N;
    const uint4* image;       // split-bf16 codebook image (prepare_codebook_tc_kernel)
    int64_t K;                // codes in the image
    int64_t k_offset;         // global index of code 0 (multiple of 256)
    int tiles_per_split;
    int l2norm;
    float t_inv;
    uint64_t seed, offset;
    int64_t row_offset;
    unsigned long long* keys; // (N) merged with atomicMax
    float* second;            // (N) deterministic modes: runner-up approximate score, atomicMax on orderable bits
};

// ---- deterministic epilogue: running top-2 of the approximate scores ---------------------------
struct DetState {
    float best, second;
    uint32_t idx;
};
__device__ __forceinline__ void det_chunk(const uint32_t (&v)[32], uint32_t code0, DetState& st) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const float m = max8(&v[8 * g]);
        if (m > st.best - kTcMargin) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float s = __uint_as_float(v[8 * g + j]);
                st.second = fmaxf(st.second, fminf(s, st.best));
                if (s > st.best) { st.best = s; st.idx = code0 + 8 * g + j; }
            }
        }
    }
}

// ---- stochastic epilogue: lazy Gumbel-max ---------------------------------------------------------
struct SampleCtx {
    float* queue;                 // this warp's ring (kQueueSlots x kQueueWords floats)
    unsigned long long* s_key;    // (128) per-row best key, shared by both epilogue groups
    uint32_t qhead, qtail;        // warp-uniform
    float vstar;                  // this lane's row: best perturbed score so far
    uint32_t row_local, grow;
    float t_inv;
    uint64_t seed, offset;
    int64_t k_end_global;         // first invalid global code
};

__device__ __forceinline__ void sample_flush(SampleCtx& c, uint32_t lane, uint32_t n) {
    if (lane < n) {
        const float* e = c.queue + ((c.qhead + lane) & (TcSmem::kQueueSlots - 1)) * TcSmem::kQueueWords;
        const float4 s0 = *reinterpret_cast<const float4*>(e);
        const float4 s1 = *reinterpret_cast<const float4*>(e + 4);
        const uint2 meta = *reinterpret_cast<const uint2*>(e + 8);
        const uint32_t group = meta.x, rl = meta.y & 0xFFu, grow = c.grow - c.row_local + rl;
        const Philox4 mw = group_max_words(group >> 2, grow, c.seed, c.offset);
        float g[8];
        group_noise(group, grow, c.seed, c.offset, pick_word(mw, group & 3u), g);
        const float s[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
        float bv = -INFINITY;
        uint32_t bk = group * 8;
#pragma unroll
        for (int p = 0; p < 8; ++p) {
            const float v = fmaf(s[p], c.t_inv, g[p]);
            if (static_cast<int64_t>(group) * 8 + p < c.k_end_global && v > bv) { bv = v; bk = group * 8 + p; }
        }
        atomicMax(c.s_key + rl, make_key(bv, bk));
    }
    c.qhead += n;
    __syncwarp();
    c.vstar = key_value(c.s_key[c.row_local]);
}

__device__ __forceinline__ void sample_chunk(const uint32_t (&v)[32], uint32_t code0_global, uint32_t lane, SampleCtx& c) {
    // code0_global is a multiple of 32: groups 4q .. 4q+3 share one Philox call
    const Philox4 mw = group_max_words(code0_global >> 5, c.grow, c.seed, c.offset);
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const float m = max8(&v[8 * g]);
        const float bound = fmaf(m, c.t_inv, group_max_noise(pick_word(mw, g)));
        const bool ev = bound > c.vstar;
        const uint32_t b = __ballot_sync(0xffffffffu, ev);
        if (b) {
            if (ev) {
                const uint32_t slot = (c.qtail + __popc(b & ((1u << lane) - 1u))) & (TcSmem::kQueueSlots - 1);
                float* e = c.queue + slot * TcSmem::kQueueWords;
                *reinterpret_cast<uint4*>(e) = make_uint4(v[8 * g], v[8 * g + 1], v[8 * g + 2], v[8 * g + 3]);
                *reinterpret_cast<uint4*>(e + 4) = make_uint4(v[8 * g + 4], v[8 * g + 5], v[8 * g + 6], v[8 * g + 7]);
                *reinterpret_cast<uint2*>(e + 8) = make_uint2((code0_global >> 3) + g, c.row_local);
            }
            c.qtail += __popc(b);
            __syncwarp();
            if (c.qtail - c.qhead >= 32u) sample_flush(c, lane, 32u);
        }
    }
}

// =======================================================================================
// grid = (row tiles of 128, code splits); block = 320 threads; dynamic smem = tc_smem_bytes(D, MODE)
// =======================================================================================
template <int D>
__host__ __device__ constexpr size_t tc_smem_bytes() {
    return 1024 /*align slack*/ + static_cast<size_t>(kTcRows) * tc_kc(D) * 2 + kTcStages * tc_tile_bytes(D) +
           8 * TcSmem::kQueueSlots * TcSmem::kQueueWords * 4 + kTcRows * 8 + 3 * kTcRows * 4 + 256;
}

template <int D, typename T, int MODE>
__global__ void __launch_bounds__(kTcThreads, 1) tc_scores_kernel(const TcParams p) {
    constexpr int KC = tc_kc(D);
    constexpr uint32_t A_BYTES = kTcRows * KC * 2;
    constexpr uint32_t B_BYTES = static_cast<uint32_t>(tc_tile_bytes(D));
    constexpr uint32_t A_LBO = kTcRows * 16, B_LBO = kTcCodes * 16, SBO = 128;
    constexpr bool SAMPLE = MODE == B200VQ_MODE_SAMPLE;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sA = smem;
    uint8_t* sB = sA + A_BYTES;
    float* s_queue = reinterpret_cast<float*>(sB + kTcStages * B_BYTES);
    unsigned long long* s_key = reinterpret_cast<unsigned long long*>(s_queue + 8 * TcSmem::kQueueSlots * TcSmem::kQueueWords);
    float* s_best = reinterpret_cast<float*>(s_key + kTcRows);
    float* s_second = s_best + kTcRows;
    uint32_t* s_idx = reinterpret_cast<uint32_t*>(s_second + kTcRows);
    uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_idx + kTcRows);  // full[S], empty[S], tfull[2], tempty[2]
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(s_bar + 2 * kTcStages + 4);

    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row0 = static_cast<int64_t>(blockIdx.x) * kTcRows;
    const int total_tiles = static_cast<int>((p.K + kTcCodes - 1) / kTcCodes);
    const int tile_begin = blockIdx.y * p.tiles_per_split;
    const int ntiles = min(p.tiles_per_split, total_tiles - tile_begin);

    const uint32_t bar_full = smem_u32(s_bar), bar_empty = bar_full + 8 * kTcStages;
    const uint32_t bar_tfull = bar_empty + 8 * kTcStages, bar_tempty = bar_tfull + 16;

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
        for (int c = 0; c < KC / 8; ++c) {
            uint32_t w[4];
#pragma unroll
            for (int e = 0; e < 4; ++e)
                w[e] = static_cast<uint32_t>(split_elem<D, true>(hi, lo, c * 8 + 2 * e)) |
                       (static_cast<uint32_t>(split_elem<D, true>(hi, lo, c * 8 + 2 * e + 1)) << 16);
            *reinterpret_cast<uint4*>(sA + c * A_LBO + threadIdx.x * 16) = make_uint4(w[0], w[1], w[2], w[3]);
        }
        s_key[threadIdx.x] = kKeyNegInf;
    } else if (warp == 8) {
        if (lane == 0) {
            for (int s = 0; s < kTcStages; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
            for (int b = 0; b < 2; ++b) { mbar_init(bar_tfull + 8 * b, 1); mbar_init(bar_tempty + 8 * b, 128); }
            fence_barrier_init();
        }
    } else if (warp == 9) {
        tmem_alloc(smem_u32(s_tmem), 512);
    }
    fence_proxy_async();  // A tile (generic-proxy stores) -> visible to the tensor core (async proxy)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;

    if (warp == 8) {
        // ===== codebook tile producer: one bulk copy per tile ========================================
        if (lane == 0) {
            const uint8_t* src = reinterpret_cast<const uint8_t*>(p.image) + static_cast<size_t>(tile_begin) * B_BYTES;
            for (int i = 0; i < ntiles; ++i) {
                const int s = i % kTcStages;
                mbar_wait(bar_empty + 8 * s, ((i / kTcStages) & 1) ^ 1);
                mbar_expect_tx(bar_full + 8 * s, B_BYTES);
                bulk_g2s(smem_u32(sB + s * B_BYTES), src + static_cast<size_t>(i) * B_BYTES, B_BYTES, bar_full + 8 * s);
            }
        }
    } else if (warp == 9) {
        // ===== MMA issuer =====================================================================================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(kTcRows, kTcCodes);
            for (int i = 0; i < ntiles; ++i) {
                const int buf = i & 1, s = i % kTcStages;
                mbar_wait(bar_tempty + 8 * buf, ((i >> 1) & 1) ^ 1);
                mbar_wait(bar_full + 8 * s, (i / kTcStages) & 1);
                tc_fence_after();
                const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB + s * B_BYTES);
#pragma unroll
                for (int k = 0; k < KC / 16; ++k) {
                    umma_bf16(tmem_base + buf * kTcCodes, umma_desc(a0 + k * 2 * A_LBO, A_LBO, SBO),
                              umma_desc(b0 + k * 2 * B_LBO, B_LBO, SBO), idesc, k > 0);
                }
                umma_commit(bar_empty + 8 * s);     // stage reusable once these MMAs have read it
                umma_commit(bar_tfull + 8 * buf);   // accumulator buffer ready for its epilogue group
            }
        }
    } else {
        // ===== epilogue groups (warps 0-3 drain buffer 0, warps 4-7 buffer 1) ==========================
        const int grp = warp >> 2;
        const uint32_t row_local = 32 * (warp & 3) + lane;
        const uint32_t t_lane = tmem_base + ((32u * (warp & 3)) << 16) + grp * kTcCodes;
        DetState det = {-INFINITY, -INFINITY, static_cast<uint32_t>(p.k_offset) + tile_begin * kTcCodes};
        SampleCtx sc;
        if (SAMPLE) {
            sc.queue = s_queue + warp * TcSmem::kQueueSlots * TcSmem::kQueueWords;
            sc.s_key = s_key;
            sc.qhead = sc.qtail = 0;
            sc.vstar = -INFINITY;
            sc.row_local = row_local;
            sc.grow = static_cast<uint32_t>(p.row_offset + row0 + row_local);
            sc.t_inv = p.t_inv;
            sc.seed = p.seed;
            sc.offset = p.offset;
            sc.k_end_global = p.k_offset + p.K;
        }
        uint32_t va[32], vb[32];
        for (int i = grp; i < ntiles; i += 2) {
            mbar_wait(bar_tfull + 8 * grp, (i >> 1) & 1);
            tc_fence_after();
            const int64_t tile_code0 = static_cast<int64_t>(tile_begin + i) * kTcCodes;  // local code index
            const int valid = static_cast<int>(min(static_cast<int64_t>(kTcCodes), p.K - tile_code0));
            tmem_ld32(t_lane, va);
#pragma unroll 1
            for (int c = 0; c < kTcCodes / 32; c += 2) {
                tmem_ld_wait32(va);
                tmem_ld32(t_lane + (c + 1) * 32, vb);
                if ((c + 1) * 32 > valid) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) if (c * 32 + j >= valid) va[j] = 0xFF800000u;
                }
                const uint32_t code_a = static_cast<uint32_t>(p.k_offset + tile_code0) + c * 32;
                if (SAMPLE) sample_chunk(va, code_a, lane, sc); else det_chunk(va, code_a, det);
                tmem_ld_wait32(vb);
                if (c + 2 < kTcCodes / 32) tmem_ld32(t_lane + (c + 2) * 32, va);
                if ((c + 2) * 32 > valid) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) if ((c + 1) * 32 + j >= valid) vb[j] = 0xFF800000u;
                }
                if (SAMPLE) sample_chunk(vb, code_a + 32, lane, sc); else det_chunk(vb, code_a + 32, det);
            }
            tc_fence_before();
            mbar_arrive(bar_tempty + 8 * grp);
        }
        if (SAMPLE) {
            while (sc.qtail != sc.qhead) sample_flush(sc, lane, min(32u, sc.qtail - sc.qhead));
        } else if (grp == 1) {
            s_best[row_local] = det.best; s_second[row_local] = det.second; s_idx[row_local] = det.idx;
        }
        // merge the two groups (named barrier over the 256 epilogue threads)
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (grp == 0) {
            const int64_t row = row0 + row_local;
            if (row < p.N && ntiles > 0) {
                if (SAMPLE) {
                    atomicMax(p.keys + row, s_key[row_local]);
                } else {
                    const float ob = s_best[row_local], os = s_second[row_local];
                    const uint32_t oi = s_idx[row_local];
                    float best = det.best, second = fmaxf(det.second, os);
                    uint32_t idx = det.idx;
                    if (ob > best || (ob == best && oi < idx)) { second = fmaxf(second, best); best = ob; idx = oi; }
                    else second = fmaxf(second, ob);
                    // cross-split merge: the loser of every key comparison is a runner-up candidate
                    const unsigned long long mine = make_key(best, idx);
                    const unsigned long long old = atomicMax(p.keys + row, mine);
                    float runner = second;
                    if (old != 0ull) runner = fmaxf(runner, fminf(key_value(old), best));
                    atomicMax(reinterpret_cast<unsigned int*>(p.second) + row, orderable(runner));
                }
            }
        }
    }
    // ---- teardown ------------------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    if (warp == 9) tmem_dealloc(tmem_base, 512);
}

}  // namespace b200vq

namespace b200vq {

// =======================================================================================
// Exact re-scoring of the rows the tensor-core pass could not decide (runner-up within
// kTcMargin of the best approximate score).  Same fp32 arithmetic as scores_kernel, so the
// final indices are bit-identical to the exact path.  Each CTA owns a contiguous range of
// rows, collects the undecided ones 256 at a time and re-scans the codebook for each with
// all 256 threads.
// =======================================================================================
template <int D, typename T, int MODE>
__global__ void __launch_bounds__(256) resolve_kernel(const T* __restrict__ z, int64_t z_row_stride, int64_t N,
                                                      const float* __restrict__ emb, const float* __restrict__ e2,
                                                      int64_t K, int64_t k_offset, int l2norm,
                                                      unsigned long long* __restrict__ keys,
                                                      const float* __restrict__ second,
                                                      unsigned int* __restrict__ n_resolved) {
    __shared__ int s_list[256];
    __shared__ int s_count;
    __shared__ unsigned long long s_red[8];
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
            if (!(best - sec >= kTcMargin)) s_list[atomicAdd(&s_count, 1)] = threadIdx.x;
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

// b200vq -- C ABI (include/b200vq.h) over the sm_100a kernels.  No torch types here.
// This unit: the tcgen05 tensor-core scoring path.
#include <type_traits>

#include "vq_tc_sample.cuh"
#include "b200vq_host.h"

using namespace b200vq;
using namespace b200vq::host;

namespace {
template <int DD, typename T, int MM>
int launch_tc(const TcParams& prm, dim3 grid, cudaStream_t st) {
    constexpr size_t smem = tc_smem_bytes<DD>();
    if (int rc = set_smem(tc_scores_kernel<DD, T, MM>, smem)) return rc;   // per device: set on every launch
    tc_scores_kernel<DD, T, MM><<<grid, kTcThreads, smem, st>>>(prm);
    return launch_check();
}
template <int DD, typename T>
int launch_ts(const TsParams& prm, dim3 grid, cudaStream_t st) {
    constexpr size_t smem = ts_smem_bytes<DD>();
    if constexpr (DD == 24 && std::is_same<T, float>::value) {
        if (prm.debug != 0) {   // tuning switches exist in this one instantiation only
            if (int rc = set_smem(tc_sample_kernel<DD, T, true>, smem)) return rc;
            tc_sample_kernel<DD, T, true><<<grid, kTsThreads, smem, st>>>(prm);
            return launch_check();
        }
    }
    if (int rc = set_smem(tc_sample_kernel<DD, T>, smem)) return rc;
    tc_sample_kernel<DD, T><<<grid, kTsThreads, smem, st>>>(prm);
    return launch_check();
}
inline int env_int(const char* name, int fallback) {
    const char* e = std::getenv(name);
    return e ? std::atoi(e) : fallback;
}
}  // namespace

namespace {
__global__ void philox_advance_kernel(unsigned long long* st, unsigned long long inc) { st[1] += inc; }

// code splits: fill the SMs, keep the last wave full, keep >= 4 tiles per CTA when possible
inline int tc_code_splits(int64_t row_tiles, int total_tiles) {
    const int sms = sm_count();
    int best_splits = 1;
    double best_cost = 1e30;
    for (int sp = 1; sp <= total_tiles; sp *= 2) {
        const int per = (total_tiles + sp - 1) / sp;
        if (sp > 1 && per < 4) break;
        const int real = (total_tiles + per - 1) / per;
        const double ctas = static_cast<double>(row_tiles) * real;
        // measured (tools/tc_probe.py): a CTA costs ~6.5 tiles of prologue/teardown on top of its tiles,
        // and CTAs of later waves start as soon as an SM frees up
        const double waves = ctas <= sms ? 1.0 : ctas / sms;
        const double cost = waves * (per + 6.5);
        if (cost < best_cost) { best_cost = cost; best_splits = sp; }
    }
    if (const char* env = std::getenv("B200VQ_TC_SPLITS")) {  // tuning aid
        const int v = std::atoi(env);
        if (v >= 1 && v <= total_tiles) best_splits = v;
    }
    return best_splits;
}
}  // namespace

namespace b200vq {
namespace host {
int launch_tc_sample(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D, const void* image, const float* emb,
                     int64_t K, int64_t k_offset, int l2norm, float t_inv, uint64_t seed, uint64_t offset,
                     unsigned long long* philox_dev, int64_t row_offset, uint64_t* keys, unsigned int* n_events, cudaStream_t st) {
    if (N < 0 || K <= 0 || k_offset < 0 || row_offset < 0 || !valid_dtype(z_dtype)) return B200VQ_EINVAL;
    if (!tc_dim_ok(D)) return B200VQ_EDIM;
    if (!l2norm) return B200VQ_EINVAL;
    if (N == 0) return B200VQ_OK;
    if (!z || !image || !emb || z_row_stride < D) return B200VQ_EINVAL;
    if (k_offset & (kTcCodes - 1)) return B200VQ_EALIGN;
    if (k_offset + K > 0xFFFFFFFFll || row_offset + N > 0xFFFFFFFFll) return B200VQ_EINVAL;
    const int64_t row_tiles = (N + kTcRows - 1) / kTcRows;
    const int total_tiles = static_cast<int>((K + kTcCodes - 1) / kTcCodes);
    const int want = tc_code_splits(row_tiles, total_tiles);
    const int per = (total_tiles + want - 1) / want;
    const int splits = (total_tiles + per - 1) / per;
    if (!keys) return B200VQ_EINVAL;
    // second-generation sampler (vq_tc_sample.cuh): bf16 filter + exact expansion of the candidates
    TsParams ts{};
    ts.z = z; ts.z_row_stride = z_row_stride; ts.N = N; ts.image = static_cast<const uint8_t*>(image); ts.emb = emb;
    ts.K = K; ts.k_offset = k_offset; ts.tiles_per_split = per; ts.l2norm = l2norm; ts.t_inv = t_inv;
    ts.seed = seed; ts.offset = offset; ts.row_offset = row_offset; ts.philox_dev = philox_dev;
    ts.keys = reinterpret_cast<unsigned long long*>(keys); ts.n_events = n_events;
    ts.debug = env_int("B200VQ_TC_DEBUG", 0);
    // warm start: scan the first eighth of the CTA's tiles for the row's best code before testing
    auto scout_for = [](int tiles) { return std::min(tiles, std::max(0, env_int("B200VQ_TS_SCOUT", tiles >= 16 ? tiles / 8 : (tiles >= 8 ? 1 : 0)))); };
    ts.scout_tiles = scout_for(per);
    ts.head_row_tiles = static_cast<int>(row_tiles); ts.splits = splits;
    ts.tail_splits = 1; ts.tail_tiles_per_split = per; ts.tail_scout_tiles = ts.scout_tiles;
    // Large batches: row_tiles / SMs waves of one CTA per SM, the last one partly filled (65 536 rows: 3.46 -> the
    // fourth wave runs 68 CTAs on 148 SMs).  Its row tiles are split over the codes so that it takes 1 / tail_splits
    // of a wave (keys of the parts merge through atomicMax like any other code split).
    const int sms = sm_count();
    const int64_t rem = row_tiles % sms;
    if (splits == 1 && row_tiles > sms && rem > 0 && env_int("B200VQ_TS_TAIL", 1) != 0) {
        int tsp = static_cast<int>(std::min<int64_t>(4, sms / rem));
        while (tsp > 1 && total_tiles / tsp < 8) --tsp;
        if (const int forced = env_int("B200VQ_TS_TAIL_SPLITS", 0)) tsp = std::max(1, std::min(forced, total_tiles));
        if (tsp > 1) {
            ts.head_row_tiles = static_cast<int>(row_tiles - rem);
            ts.tail_splits = tsp;
            ts.tail_tiles_per_split = (total_tiles + tsp - 1) / tsp;
            ts.tail_scout_tiles = scout_for(ts.tail_tiles_per_split);
        }
    }
    ts.flush_early = std::max(1, env_int("B200VQ_TS_FLUSH_EARLY", 8));
    ts.flush_late = std::max(1, env_int("B200VQ_TS_FLUSH_LATE", 32));
    ts.early_iters = env_int("B200VQ_TS_EARLY_ITERS", 2);
    dim3 grid(static_cast<unsigned>(static_cast<int64_t>(ts.head_row_tiles) * splits + (row_tiles - ts.head_row_tiles) * ts.tail_splits));
    int rc = B200VQ_OK;
    B200VQ_DISPATCH_DTYPE(z_dtype, T, {
        switch (D) {
            case 8: rc = launch_ts<8, T>(ts, grid, st); break;
            case 16: rc = launch_ts<16, T>(ts, grid, st); break;
            case 24: rc = launch_ts<24, T>(ts, grid, st); break;
            default: rc = launch_ts<32, T>(ts, grid, st); break;
        }
    });
    if (rc) return rc;
    if (philox_dev) {
        philox_advance_kernel<<<1, 1, 0, st>>>(philox_dev, 4ull);
        rc = launch_check();
    }
    return rc;
}
}  // namespace host
}  // namespace b200vq

extern "C" {

// ---- tensor-core path --------------------------------------------------------------------------
size_t b200vq_tc_scratch_bytes(int64_t N) {
    // [runner-up scores: N f32][undecided-row list: count + 3 pad, N u32]; the caller zeroes it
    return N < 0 ? 0 : static_cast<size_t>(N) * 8 + 16;
}

int b200vq_tc_supported(int D, int l2norm, int mode) {
    return tc_dim_ok(D) && l2norm && valid_mode(mode) ? 1 : 0;
}

size_t b200vq_tc_image_bytes(int64_t K, int D) {
    if (K <= 0 || !tc_dim_ok(D)) return 0;
    // tiles, then one 16-byte trailer holding the "codebook has a non-unit row" flag word
    return static_cast<size_t>((K + kTcCodes - 1) / kTcCodes) * tc_tile_bytes(D) + 16;
}

int b200vq_prepare_codebook_tc(const float* emb, int64_t K, int D, void* image, b200vq_stream_t stream) {
    if (!emb || !image || K <= 0) return B200VQ_EINVAL;
    if (!tc_dim_ok(D)) return B200VQ_EDIM;
    if (reinterpret_cast<uintptr_t>(image) & 15) return B200VQ_EALIGN;
    const unsigned blocks = static_cast<unsigned>((K + kTcCodes - 1) / kTcCodes);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint4* img = static_cast<uint4*>(image);
    unsigned int* flags = reinterpret_cast<unsigned int*>(static_cast<char*>(image) + static_cast<size_t>(blocks) * tc_tile_bytes(D));
    int rc = cuda_check(cudaMemsetAsync(flags, 0, 16, st));
    if (rc) return rc;
    switch (D) {
        case 8: prepare_codebook_tc_kernel<8><<<blocks, kTcCodes, 0, st>>>(emb, K, img, flags); break;
        case 16: prepare_codebook_tc_kernel<16><<<blocks, kTcCodes, 0, st>>>(emb, K, img, flags); break;
        case 24: prepare_codebook_tc_kernel<24><<<blocks, kTcCodes, 0, st>>>(emb, K, img, flags); break;
        default: prepare_codebook_tc_kernel<32><<<blocks, kTcCodes, 0, st>>>(emb, K, img, flags); break;
    }
    return launch_check();
}


int b200vq_scores_tc(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D, const void* image,
                     const float* emb, const float* e2, int64_t K, int64_t k_offset, int mode, int l2norm,
                     float t_inv, uint64_t seed, uint64_t offset, int64_t row_offset, uint64_t* keys, float* scratch,
                     unsigned int* n_resolved, b200vq_stream_t stream) {
    if (N < 0 || K <= 0 || k_offset < 0 || row_offset < 0 || !valid_dtype(z_dtype) || !valid_mode(mode)) return B200VQ_EINVAL;
    if (!tc_dim_ok(D)) return B200VQ_EDIM;
    if (!l2norm) return B200VQ_EINVAL;  // the error bound assumes unit vectors
    if (N == 0) return B200VQ_OK;
    if (!z || !image || !keys || z_row_stride < D) return B200VQ_EINVAL;
    if (!emb || (mode != B200VQ_MODE_SAMPLE && (!e2 || !scratch))) return B200VQ_EINVAL;
    if (k_offset & (kTcCodes - 1)) return B200VQ_EALIGN;
    if (k_offset + K > 0xFFFFFFFFll || row_offset + N > 0xFFFFFFFFll) return B200VQ_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t row_tiles = (N + kTcRows - 1) / kTcRows;
    const int total_tiles = static_cast<int>((K + kTcCodes - 1) / kTcCodes);
    const int best_splits = tc_code_splits(row_tiles, total_tiles);
    const int per = (total_tiles + best_splits - 1) / best_splits;
    const int splits = (total_tiles + per - 1) / per;
    TcParams prm;
    prm.z = z; prm.z_row_stride = z_row_stride; prm.N = N; prm.image = static_cast<const uint4*>(image);
    prm.K = K; prm.k_offset = k_offset; prm.tiles_per_split = per; prm.l2norm = l2norm; prm.t_inv = t_inv;
    prm.seed = seed; prm.offset = offset; prm.row_offset = row_offset;
    prm.keys = reinterpret_cast<unsigned long long*>(keys); prm.second = scratch;
    prm.debug = 0;
    prm.n_events = mode == B200VQ_MODE_SAMPLE ? n_resolved : nullptr;
    if (const char* env = std::getenv("B200VQ_TC_DEBUG")) prm.debug = std::atoi(env);
    dim3 grid(static_cast<unsigned>(row_tiles), static_cast<unsigned>(splits));
    int rc = B200VQ_OK;
    if (mode == B200VQ_MODE_SAMPLE)
        return launch_tc_sample(z, z_dtype, z_row_stride, N, D, image, emb, K, k_offset, l2norm, t_inv, seed, offset, nullptr,
                                row_offset, keys, n_resolved, st);
#define B200VQ_TC_DET(DD)                                                                         \
    if (mode == B200VQ_MODE_ARGMIN_DIST) rc = launch_tc<DD, T, B200VQ_MODE_ARGMIN_DIST>(prm, grid, st); \
    else rc = launch_tc<DD, T, B200VQ_MODE_ARGMAX_COS>(prm, grid, st)
    B200VQ_DISPATCH_DTYPE(z_dtype, T, {
        switch (D) {
            case 8: B200VQ_TC_DET(8); break;
            case 16: B200VQ_TC_DET(16); break;
            case 24: B200VQ_TC_DET(24); break;
            default: B200VQ_TC_DET(32); break;
        }
    });
#undef B200VQ_TC_DET
    if (rc || prm.debug) return rc;
    // exact re-scoring of undecided rows (deterministic modes)
    const unsigned int* flags = reinterpret_cast<const unsigned int*>(
        static_cast<const char*>(image) + static_cast<size_t>(total_tiles) * tc_tile_bytes(D));
    unsigned long long* k64 = reinterpret_cast<unsigned long long*>(keys);
    unsigned int* list = reinterpret_cast<unsigned int*>(scratch + N);
#define B200VQ_GROUPS(DD, MM)                                                                                  \
    resolve_groups_kernel<DD, T, MM><<<static_cast<unsigned>((N + kResolveRows - 1) / kResolveRows), 256, 0, st>>>(                      \
        static_cast<const T*>(z), z_row_stride, N, emb, e2, K, k_offset, l2norm, k64, scratch, flags, list)
    B200VQ_DISPATCH_DTYPE(z_dtype, T, {
        if (mode == B200VQ_MODE_ARGMIN_DIST) {
            switch (D) {
                case 8: B200VQ_GROUPS(8, B200VQ_MODE_ARGMIN_DIST); break;
                case 16: B200VQ_GROUPS(16, B200VQ_MODE_ARGMIN_DIST); break;
                case 24: B200VQ_GROUPS(24, B200VQ_MODE_ARGMIN_DIST); break;
                default: B200VQ_GROUPS(32, B200VQ_MODE_ARGMIN_DIST); break;
            }
        } else {
            switch (D) {
                case 8: B200VQ_GROUPS(8, B200VQ_MODE_ARGMAX_COS); break;
                case 16: B200VQ_GROUPS(16, B200VQ_MODE_ARGMAX_COS); break;
                case 24: B200VQ_GROUPS(24, B200VQ_MODE_ARGMAX_COS); break;
                default: B200VQ_GROUPS(32, B200VQ_MODE_ARGMAX_COS); break;
            }
        }
    });
#undef B200VQ_GROUPS
    rc = launch_check();
    if (rc) return rc;
    const unsigned rblocks = static_cast<unsigned>(kRescoreSegs * std::min<int64_t>(N, 2ll * sm_count()));
#define B200VQ_RESOLVE(DD, MM)                                                                                \
    rescore_kernel<DD, T, MM><<<rblocks, 256, 0, st>>>(static_cast<const T*>(z), z_row_stride, emb, e2, K,    \
                                                        k_offset, l2norm, k64, list, n_resolved)
    B200VQ_DISPATCH_DTYPE(z_dtype, T, {
        if (mode == B200VQ_MODE_ARGMIN_DIST) {
            switch (D) {
                case 8: B200VQ_RESOLVE(8, B200VQ_MODE_ARGMIN_DIST); break;
                case 16: B200VQ_RESOLVE(16, B200VQ_MODE_ARGMIN_DIST); break;
                case 24: B200VQ_RESOLVE(24, B200VQ_MODE_ARGMIN_DIST); break;
                default: B200VQ_RESOLVE(32, B200VQ_MODE_ARGMIN_DIST); break;
            }
        } else {
            switch (D) {
                case 8: B200VQ_RESOLVE(8, B200VQ_MODE_ARGMAX_COS); break;
                case 16: B200VQ_RESOLVE(16, B200VQ_MODE_ARGMAX_COS); break;
                case 24: B200VQ_RESOLVE(24, B200VQ_MODE_ARGMAX_COS); break;
                default: B200VQ_RESOLVE(32, B200VQ_MODE_ARGMAX_COS); break;
            }
        }
    });
#undef B200VQ_RESOLVE
    return launch_check();
}

}  // extern "C"

// b200vq -- C ABI (include/b200vq.h) over the sm_100a kernels.  No torch types here.
#include "vq_generic.cuh"
#include "vq_tc_kernel.cuh"
#include "vq_linear.cuh"
#include "vq_gradout.cuh"
#include "vq_entropy.cuh"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>

using namespace b200vq;

namespace {

thread_local int g_last_cuda_error = 0;

inline int cuda_check(cudaError_t e) {
    if (e == cudaSuccess) return B200VQ_OK;
    g_last_cuda_error = static_cast<int>(e);
    cudaGetLastError();  // clear the sticky launch error so later calls can proceed
    return B200VQ_ECUDA;
}
std::atomic<unsigned long long> g_launches{0};
inline int launch_check() { g_launches.fetch_add(1, std::memory_order_relaxed); return cuda_check(cudaPeekAtLastError()); }

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

inline bool valid_dtype(int dt) { return dt == B200VQ_DT_F32 || dt == B200VQ_DT_BF16 || dt == B200VQ_DT_F16; }
inline bool valid_mode(int m) { return m >= B200VQ_MODE_ARGMIN_DIST && m <= B200VQ_MODE_SAMPLE; }
inline bool specialised_dim(int D) { return D == 8 || D == 16 || D == 24 || D == 32 || D == 64; }

constexpr size_t kAlign = 256;
inline size_t align_up(size_t v) { return (v + kAlign - 1) / kAlign * kAlign; }

// ---- dispatch helpers -------------------------------------------------------------------
#define B200VQ_DISPATCH_DTYPE(dt, T, ...)                                              \
    switch (dt) {                                                                      \
        case B200VQ_DT_F32: { using T = float; __VA_ARGS__; break; }                   \
        case B200VQ_DT_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; }          \
        default: { using T = __half; __VA_ARGS__; break; }                             \
    }
#define B200VQ_DISPATCH_DIM(D, DD, ...)                                                \
    switch (D) {                                                                       \
        case 8: { constexpr int DD = 8; __VA_ARGS__; break; }                          \
        case 16: { constexpr int DD = 16; __VA_ARGS__; break; }                        \
        case 24: { constexpr int DD = 24; __VA_ARGS__; break; }                        \
        case 32: { constexpr int DD = 32; __VA_ARGS__; break; }                        \
        default: { constexpr int DD = 64; __VA_ARGS__; break; }                        \
    }
#define B200VQ_DISPATCH_MODE(m, MM, ...)                                               \
    switch (m) {                                                                       \
        case B200VQ_MODE_ARGMIN_DIST: { constexpr int MM = B200VQ_MODE_ARGMIN_DIST; __VA_ARGS__; break; } \
        case B200VQ_MODE_ARGMAX_COS: { constexpr int MM = B200VQ_MODE_ARGMAX_COS; __VA_ARGS__; break; }   \
        default: { constexpr int MM = B200VQ_MODE_SAMPLE; __VA_ARGS__; break; }        \
    }

}  // namespace

namespace {
template <int DD, typename T, int MM>
int launch_tc(const TcParams& prm, dim3 grid, cudaStream_t st) {
    constexpr size_t smem = tc_smem_bytes<DD>();
    static bool configured = false;  // per instantiation
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(tc_scores_kernel<DD, T, MM>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(smem));
        if (e != cudaSuccess) return cuda_check(e);
        configured = true;
    }
    tc_scores_kernel<DD, T, MM><<<grid, kTcThreads, smem, st>>>(prm);
    return launch_check();
}
}  // namespace

extern "C" {

int b200vq_version(void) { return B200VQ_VERSION; }
int b200vq_last_cuda_error(void) { return g_last_cuda_error; }
int b200vq_max_dim(void) { return kMaxDim; }
uint64_t b200vq_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

const char* b200vq_error_string(int code) {
    switch (code) {
        case B200VQ_OK: return "ok";
        case B200VQ_EINVAL: return "invalid argument (shape, size or enum)";
        case B200VQ_EDIM: return "unsupported code dimension";
        case B200VQ_EALIGN: return "pointer or offset not aligned as required";
        case B200VQ_ECUDA: return "CUDA error (see b200vq_last_cuda_error)";
        case B200VQ_EWORKSPACE: return "workspace null or too small";
        default: return "unknown error";
    }
}

size_t b200vq_workspace_bytes(int64_t N, int D, int64_t K) {
    (void)D; (void)K;
    if (N < 0) return 0;
    // [keys: N u64][sum + ticket, 2 doubles][resolved-row counter][tensor-core scratch]
    return align_up(static_cast<size_t>(N) * sizeof(uint64_t)) + align_up(2 * sizeof(double)) +
           align_up(sizeof(unsigned int)) + align_up(b200vq_tc_scratch_bytes(N));
}

int b200vq_prepare_codebook(const float* weight, int64_t K, int D, int l2norm, float* emb, float* e2,
                            float* wnorm, b200vq_stream_t stream) {
    if (!weight || !emb || !e2 || !wnorm || K <= 0 || D <= 0 || D > kMaxDim) return B200VQ_EINVAL;
    const int64_t threads = K * 32;
    const unsigned blocks = static_cast<unsigned>((threads + 255) / 256);
    prepare_codebook_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(weight, K, D, l2norm, emb, e2, wnorm);
    return launch_check();
}

int b200vq_keys_reset(uint64_t* keys, int64_t N, b200vq_stream_t stream) {
    if (N < 0 || (N > 0 && !keys)) return B200VQ_EINVAL;
    if (N == 0) return B200VQ_OK;
    return cuda_check(cudaMemsetAsync(keys, 0, static_cast<size_t>(N) * sizeof(uint64_t), static_cast<cudaStream_t>(stream)));
}

int b200vq_scores_partial(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D, const float* emb,
                          const float* e2, int64_t K_local, int64_t k_offset, int mode, int l2norm, float t_inv,
                          uint64_t seed, uint64_t offset, int64_t row_offset, uint64_t* keys,
                          b200vq_stream_t stream) {
    if (N < 0 || K_local < 0 || k_offset < 0 || row_offset < 0 || !valid_dtype(z_dtype) || !valid_mode(mode))
        return B200VQ_EINVAL;
    if (D <= 0 || D > kMaxDim) return B200VQ_EDIM;
    if (N == 0 || K_local == 0) return B200VQ_OK;
    if (!z || !emb || !e2 || !keys || z_row_stride < D) return B200VQ_EINVAL;
    if (k_offset + K_local > 0xFFFFFFFFll || row_offset + N > 0xFFFFFFFFll) return B200VQ_EINVAL;
    if (mode == B200VQ_MODE_SAMPLE && (k_offset & 7)) return B200VQ_EALIGN;  // noise groups of 8 codes
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned long long* k64 = reinterpret_cast<unsigned long long*>(keys);
    if (specialised_dim(D)) {
        const int64_t row_tiles = (N + 127) / 128;
        B200VQ_DISPATCH_DIM(D, DD, {
            constexpr int TK = TileCodes<DD>::value;
            // enough CTAs for >= 4 waves-worth of residency when the token count is small
            const int64_t want = 4ll * sm_count();
            int64_t splits = std::max<int64_t>(1, (want + row_tiles - 1) / row_tiles);
            const int64_t max_splits = (K_local + TK - 1) / TK;
            splits = std::min(splits, max_splits);
            int64_t per = (K_local + splits - 1) / splits;
            per = (per + TK - 1) / TK * TK;
            splits = (K_local + per - 1) / per;
            dim3 grid(static_cast<unsigned>(row_tiles), static_cast<unsigned>(splits));
            B200VQ_DISPATCH_DTYPE(z_dtype, T, {
                B200VQ_DISPATCH_MODE(mode, MM, {
                    scores_kernel<DD, T, MM><<<grid, 128, 0, st>>>(static_cast<const T*>(z), z_row_stride, N, emb, e2,
                                                                    K_local, k_offset, per, l2norm, t_inv, seed,
                                                                    offset, row_offset, k64);
                });
            });
        });
    } else {
        const int rows = D <= 512 ? 64 : 32;   // rows per CTA: [rows][D+1] floats of shared memory (<= 131 KB)
        const int64_t row_tiles = (N + rows - 1) / rows;
        const int64_t want = 8ll * sm_count();
        int64_t splits = std::max<int64_t>(1, (want + row_tiles - 1) / row_tiles);
        splits = std::min<int64_t>(splits, (K_local + 63) / 64);
        int64_t per = (K_local + splits - 1) / splits;
        per = (per + 7) / 8 * 8;
        splits = (K_local + per - 1) / per;
        dim3 grid(static_cast<unsigned>(row_tiles), static_cast<unsigned>(splits));
        const size_t smem = static_cast<size_t>(rows) * (D + 1) * sizeof(float);
        B200VQ_DISPATCH_DTYPE(z_dtype, T, {
            B200VQ_DISPATCH_MODE(mode, MM, {
                if (smem > 48 * 1024)
                    cudaFuncSetAttribute(scores_generic_kernel<T, MM>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(smem));
                scores_generic_kernel<T, MM><<<grid, rows, smem, st>>>(static_cast<const T*>(z), z_row_stride, N, D, emb,
                                                                     e2, K_local, k_offset, per, l2norm, t_inv, seed,
                                                                     offset, row_offset, k64);
            });
        });
    }
    return launch_check();
}

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
    if (mode != B200VQ_MODE_SAMPLE && (!emb || !e2 || !scratch)) return B200VQ_EINVAL;
    if (k_offset & (kTcCodes - 1)) return B200VQ_EALIGN;
    if (k_offset + K > 0xFFFFFFFFll || row_offset + N > 0xFFFFFFFFll) return B200VQ_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t row_tiles = (N + kTcRows - 1) / kTcRows;
    const int total_tiles = static_cast<int>((K + kTcCodes - 1) / kTcCodes);
    // code splits: fill the SMs, keep the last wave full, keep >= 4 tiles per CTA when possible
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
    B200VQ_DISPATCH_DTYPE(z_dtype, T, {
        switch (D) {
            case 8: { B200VQ_DISPATCH_MODE(mode, MM, { rc = launch_tc<8, T, MM>(prm, grid, st); }); break; }
            case 16: { B200VQ_DISPATCH_MODE(mode, MM, { rc = launch_tc<16, T, MM>(prm, grid, st); }); break; }
            case 24: { B200VQ_DISPATCH_MODE(mode, MM, { rc = launch_tc<24, T, MM>(prm, grid, st); }); break; }
            default: { B200VQ_DISPATCH_MODE(mode, MM, { rc = launch_tc<32, T, MM>(prm, grid, st); }); break; }
        }
    });
    if (rc || mode == B200VQ_MODE_SAMPLE || prm.debug) return rc;
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
    const unsigned rblocks = static_cast<unsigned>(kRescoreSegs * std::min<int64_t>(N, 2ll * sms));
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

int b200vq_finish(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D, const float* emb, int64_t K,
                  const uint64_t* keys, const float* q_rows, int l2norm, int64_t n_first, float* znorm_out, float* ste_out,
                  int64_t* idx_out, float* zlen_out, int32_t* hist, int32_t* hist_first, double* sqerr, int finalize,
                  int64_t N_total, float beta, float w_cb, float* losses, b200vq_stream_t stream) {
    if (N < 0 || K <= 0 || !valid_dtype(z_dtype)) return B200VQ_EINVAL;
    if (D <= 0 || D > kMaxDim) return B200VQ_EDIM;
    if (N == 0) return B200VQ_OK;
    if (!z || (!emb && !q_rows) || !keys || !znorm_out || !ste_out || !idx_out || !zlen_out || !sqerr || z_row_stride < D)
        return B200VQ_EINVAL;
    if (finalize && (!losses || N_total <= 0)) return B200VQ_EINVAL;
    if (specialised_dim(D) && ((reinterpret_cast<uintptr_t>(znorm_out) | reinterpret_cast<uintptr_t>(ste_out) |
                                reinterpret_cast<uintptr_t>(emb) | reinterpret_cast<uintptr_t>(q_rows)) & 15))
        return B200VQ_EALIGN;   // the dense fp32 tensors are accessed 16 bytes at a time
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned blocks = static_cast<unsigned>((N + 127) / 128);
    const unsigned long long* k64 = reinterpret_cast<const unsigned long long*>(keys);
    if (specialised_dim(D)) {
        B200VQ_DISPATCH_DIM(D, DD, {
            B200VQ_DISPATCH_DTYPE(z_dtype, T, {
                finish_kernel<DD, T><<<blocks, 128, 0, st>>>(static_cast<const T*>(z), z_row_stride, N, emb, K, k64,
                                                             q_rows, l2norm, n_first, znorm_out, ste_out, idx_out, zlen_out,
                                                             hist, hist_first, sqerr, finalize, N_total, beta, w_cb,
                                                             losses);
            });
        });
    } else {
        B200VQ_DISPATCH_DTYPE(z_dtype, T, {
            finish_generic_kernel<T><<<blocks, 128, 0, st>>>(static_cast<const T*>(z), z_row_stride, N, D, emb, K, k64,
                                                             q_rows, l2norm, n_first, znorm_out, ste_out, idx_out, zlen_out,
                                                             hist, hist_first, sqerr, finalize, N_total, beta, w_cb,
                                                             losses);
        });
    }
    return launch_check();
}

int b200vq_forward(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D, const float* emb,
                   const float* e2, const void* image, int64_t K, int mode, int l2norm, float t_inv, uint64_t seed, uint64_t offset,
                   int64_t n_first, float beta, float w_cb, float* znorm_out, float* ste_out, int64_t* idx_out,
                   float* zlen_out, int32_t* hist, int32_t* hist_first, float* losses, void* workspace,
                   size_t workspace_bytes, b200vq_stream_t stream) {
    if (N <= 0 || K <= 0) return B200VQ_EINVAL;
    if (!workspace || workspace_bytes < b200vq_workspace_bytes(N, D, K)) return B200VQ_EWORKSPACE;
    if (reinterpret_cast<uintptr_t>(workspace) & 15) return B200VQ_EALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint64_t* keys = static_cast<uint64_t*>(workspace);
    double* scratch = reinterpret_cast<double*>(static_cast<char*>(workspace) + align_up(static_cast<size_t>(N) * 8));
    int rc = cuda_check(cudaMemsetAsync(workspace, 0, b200vq_workspace_bytes(N, D, K), st));
    if (rc) return rc;
    if (hist && (rc = cuda_check(cudaMemsetAsync(hist, 0, static_cast<size_t>(K) * 4, st)))) return rc;
    if (hist_first && (rc = cuda_check(cudaMemsetAsync(hist_first, 0, static_cast<size_t>(K) * 4, st)))) return rc;
    if (image && b200vq_tc_supported(D, l2norm, mode)) {
        char* base = static_cast<char*>(workspace) + align_up(static_cast<size_t>(N) * 8) + align_up(2 * sizeof(double));
        unsigned int* n_resolved = reinterpret_cast<unsigned int*>(base);
        float* second = reinterpret_cast<float*>(base + align_up(sizeof(unsigned int)));
        rc = b200vq_scores_tc(z, z_dtype, z_row_stride, N, D, image, emb, e2, K, 0, mode, l2norm, t_inv, seed, offset, 0,
                              keys, second, n_resolved, stream);
    } else {
        rc = b200vq_scores_partial(z, z_dtype, z_row_stride, N, D, emb, e2, K, 0, mode, l2norm, t_inv, seed, offset, 0,
                                   keys, stream);
    }
    if (rc) return rc;
    return b200vq_finish(z, z_dtype, z_row_stride, N, D, emb, K, keys, nullptr, l2norm, n_first, znorm_out, ste_out, idx_out,
                         zlen_out, hist, hist_first, scratch, 1, N, beta, w_cb, losses, stream);
}

int b200vq_backward(const float* g_ste, const float* g_znorm, const float* g_loss, const float* znorm,
                    const float* zlen, const int64_t* idx, const float* emb, const float* q_rows,
                    int64_t k_offset, int64_t N, int D, int64_t K, int l2norm, int64_t N_total, float beta, float w_cb, void* grad_z, int gz_dtype, float* grad_emb,
                    b200vq_stream_t stream) {
    if (N < 0 || K <= 0 || N_total <= 0 || !valid_dtype(gz_dtype)) return B200VQ_EINVAL;
    if (D <= 0 || D > kMaxDim) return B200VQ_EDIM;
    if (!grad_emb) return B200VQ_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc = cuda_check(cudaMemsetAsync(grad_emb, 0, static_cast<size_t>(K) * D * sizeof(float), st));
    if (rc) return rc;
    if (N == 0) return B200VQ_OK;
    if (!znorm || !zlen || !idx || (!emb && !q_rows) || !grad_z || k_offset < 0) return B200VQ_EINVAL;
    if (specialised_dim(D) && ((reinterpret_cast<uintptr_t>(znorm) | reinterpret_cast<uintptr_t>(g_ste) |
                                reinterpret_cast<uintptr_t>(g_znorm) | reinterpret_cast<uintptr_t>(emb) |
                                reinterpret_cast<uintptr_t>(q_rows) | reinterpret_cast<uintptr_t>(grad_z)) & 15))
        return B200VQ_EALIGN;   // the dense tensors are accessed 16 bytes at a time
    const unsigned blocks = static_cast<unsigned>((N + 127) / 128);
    if (specialised_dim(D)) {
        B200VQ_DISPATCH_DIM(D, DD, {
            B200VQ_DISPATCH_DTYPE(gz_dtype, T, {
                backward_kernel<DD, T><<<blocks, 128, 0, st>>>(g_ste, g_znorm, g_loss, znorm, zlen, idx, emb, q_rows, k_offset, K, N, l2norm,
                                                               N_total, beta, w_cb, static_cast<T*>(grad_z), grad_emb);
            });
        });
    } else {
        B200VQ_DISPATCH_DTYPE(gz_dtype, T, {
            backward_generic_kernel<T><<<blocks, 128, 0, st>>>(g_ste, g_znorm, g_loss, znorm, zlen, idx, emb, q_rows, k_offset, K, N, D,
                                                               l2norm, N_total, beta, w_cb, static_cast<T*>(grad_z),
                                                               grad_emb);
        });
    }
    return launch_check();
}

int b200vq_codebook_backward(const float* grad_emb, const float* g_emb, const float* emb, const float* wnorm,
                             int64_t K, int D, int l2norm, float* grad_weight, b200vq_stream_t stream) {
    if (!grad_emb || !emb || !wnorm || !grad_weight || K <= 0 || D <= 0 || D > kMaxDim) return B200VQ_EINVAL;
    const unsigned blocks = static_cast<unsigned>((K * 32 + 255) / 256);
    codebook_backward_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(grad_emb, g_emb, emb, wnorm, K, D,
                                                                                   l2norm, grad_weight);
    return launch_check();
}

int b200vq_decode(const int64_t* idx, int64_t n, const float* weight, int64_t K, int D, int l2norm, float* out,
                  int* bad_flag, b200vq_stream_t stream) {
    if (n < 0 || K <= 0 || D <= 0 || D > kMaxDim) return B200VQ_EINVAL;
    if (n == 0) return B200VQ_OK;
    if (!idx || !weight || !out) return B200VQ_EINVAL;
    const unsigned blocks = static_cast<unsigned>((n * 32 + 255) / 256);
    decode_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(idx, n, weight, K, D, l2norm, out, bad_flag);
    return launch_check();
}

int b200vq_usage_stats(const int32_t* hist_first, const int32_t* hist_batch, int64_t K, float* out4,
                       b200vq_stream_t stream) {
    if (!hist_first || !hist_batch || !out4 || K <= 0) return B200VQ_EINVAL;
    usage_stats_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(hist_first, hist_batch, K, out4);
    return launch_check();
}

int b200vq_gather_owned(const uint64_t* keys, int64_t N, const float* emb_local, int64_t K_local, int64_t k_offset,
                        int D, float* out, b200vq_stream_t stream) {
    if (N < 0 || K_local < 0 || D <= 0) return B200VQ_EINVAL;
    if (N == 0) return B200VQ_OK;
    if (!keys || !out || (K_local > 0 && !emb_local)) return B200VQ_EINVAL;
    const int64_t total = N * D;
    gather_owned_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const unsigned long long*>(keys), N, emb_local, K_local, k_offset, D, out);
    return launch_check();
}

}  // extern "C"

// ---- the Bottleneck's skinny linears ----------------------------------------------------------------
namespace {
inline bool lin_dim_ok(int D) { return D == 8 || D == 16 || D == 24 || D == 32; }
template <typename K>
inline int set_smem(K kernel, size_t bytes) {
    if (bytes <= 48 * 1024) return B200VQ_OK;
    return cuda_check(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)));
}
}  // namespace

extern "C" {

int b200vq_linear_supported(int C, int D) {
    if (!lin_dim_ok(D) || C < 16 || (C % 16) != 0) return 0;
    const size_t in_smem = static_cast<size_t>(D) * (C + 16) * 4, out_smem = static_cast<size_t>(C) * (D + 5) * 4;
    return in_smem <= 200 * 1024 && out_smem <= 200 * 1024 ? 1 : 0;
}

namespace {
int launch_linear_in(const float* x, int64_t ldx, int64_t N, int C, const float* w, int w_is_cd, const float* bias, int D,
                     float* y, bool precise, cudaStream_t st) {
    if (N < 0 || !b200vq_linear_supported(C, D)) return N < 0 ? B200VQ_EINVAL : B200VQ_EDIM;
    if (N == 0) return B200VQ_OK;
    if (!x || !w || !y || ldx < C) return B200VQ_EINVAL;
    if ((ldx & 3) || (reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(w) & 15) ||
        (reinterpret_cast<uintptr_t>(y) & 7))
        return B200VQ_EALIGN;
    const size_t smem = static_cast<size_t>(D) * (C + 16) * 4;
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((N + kLinInRows - 1) / kLinInRows, static_cast<int64_t>(sm_count())));
    int rc = B200VQ_OK;
#define B200VQ_LIN_IN(DD)                                                                        \
    if (precise) {                                                                               \
        if ((rc = set_smem(linear_in_kernel<DD, true>, smem))) return rc;                        \
        linear_in_kernel<DD, true><<<blocks, kLinInThreads, smem, st>>>(x, ldx, N, C, w, w_is_cd, bias, y);  \
    } else {                                                                                     \
        if ((rc = set_smem(linear_in_kernel<DD, false>, smem))) return rc;                       \
        linear_in_kernel<DD, false><<<blocks, kLinInThreads, smem, st>>>(x, ldx, N, C, w, w_is_cd, bias, y); \
    }
    switch (D) {
        case 8: B200VQ_LIN_IN(8); break;
        case 16: B200VQ_LIN_IN(16); break;
        case 24: B200VQ_LIN_IN(24); break;
        default: B200VQ_LIN_IN(32); break;
    }
#undef B200VQ_LIN_IN
    return launch_check();
}
}  // namespace

int b200vq_linear_in(const float* x, int64_t ldx, int64_t N, int C, const float* w, const float* bias, int D, float* y,
                     b200vq_stream_t stream) {
    return launch_linear_in(x, ldx, N, C, w, 0, bias, D, y, true, static_cast<cudaStream_t>(stream));
}

namespace {
int launch_linear_out(const float* a, int64_t N, int D, const float* w, int w_is_dc, const float* bias, int C, float* y,
                      int64_t ldy, const int64_t* gidx, int64_t gK, int gl2, const float* add, int64_t add_rows, int* bad_flag,
                      cudaStream_t st) {
    if (N < 0 || !b200vq_linear_supported(C, D)) return N < 0 ? B200VQ_EINVAL : B200VQ_EDIM;
    if (N == 0) return B200VQ_OK;
    if (!a || !w || !y || ldy < C) return B200VQ_EINVAL;
    if ((ldy & 3) || (reinterpret_cast<uintptr_t>(y) & 15) || (reinterpret_cast<uintptr_t>(w) & 15)) return B200VQ_EALIGN;
    if (add && (add_rows <= 0 || (reinterpret_cast<uintptr_t>(add) & 15))) return add_rows <= 0 ? B200VQ_EINVAL : B200VQ_EALIGN;
    const size_t smem = static_cast<size_t>(C) * (D + 5) * 4;
    int lo_dbg = 0, lo_grid = 1;   // one 16-warp CTA per SM: the weight staging is paid once per SM
    if (const char* env = std::getenv("B200VQ_LO_DEBUG")) lo_dbg = std::atoi(env);
    if (const char* env = std::getenv("B200VQ_LO_GRID")) lo_grid = std::max(1, std::atoi(env));
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((N + kLinRows - 1) / kLinRows, static_cast<int64_t>(lo_grid) * sm_count()));
    int rc = B200VQ_OK;
#define B200VQ_LIN_OUT(DD)                                                           \
    if ((rc = set_smem(linear_out_kernel<DD>, smem))) return rc;                     \
    linear_out_kernel<DD><<<blocks, kLinOutThreads, smem, st>>>(a, N, w, w_is_dc, bias, C, y, ldy, lo_dbg, gidx, gK, gl2, add, \
                                                                add_rows, bad_flag)
    switch (D) {
        case 8: B200VQ_LIN_OUT(8); break;
        case 16: B200VQ_LIN_OUT(16); break;
        case 24: B200VQ_LIN_OUT(24); break;
        default: B200VQ_LIN_OUT(32); break;
    }
#undef B200VQ_LIN_OUT
    return launch_check();
}
}  // namespace

int b200vq_linear_out(const float* a, int64_t N, int D, const float* w, int w_is_dc, const float* bias, int C, float* y,
                      int64_t ldy, b200vq_stream_t stream) {
    return launch_linear_out(a, N, D, w, w_is_dc, bias, C, y, ldy, nullptr, 0, 0, nullptr, 0, nullptr,
                             static_cast<cudaStream_t>(stream));
}

int b200vq_decode_project(const int64_t* idx, int64_t n, const float* weight, int64_t K, int D, int l2norm, const float* w_out,
                          const float* b_out, int C, const float* add, int64_t add_rows, float* out, int64_t ldo, int* bad_flag,
                          b200vq_stream_t stream) {
    if (n < 0 || K <= 0) return B200VQ_EINVAL;
    if (n > 0 && !idx) return B200VQ_EINVAL;
    return launch_linear_out(weight, n, D, w_out, 0, b_out, C, out, ldo, idx, K, l2norm, add, add_rows, bad_flag,
                             static_cast<cudaStream_t>(stream));
}

extern "C++" {
namespace {
template <int D, bool WITH_GA>
int launch_grad_out_d(const float* g, int64_t ldg, const float* a, const float* w, int64_t N, int C, const float* ga_add, float* ga,
                      float* partial, unsigned blocks, cudaStream_t st) {
    const size_t smem = WITH_GA ? go_smem_bytes<D>(C) : go_smem_bytes_wgrad<D>();
    int rc = B200VQ_OK;
#define B200VQ_GO(NS)                                                                       \
    if ((rc = set_smem(grad_out_kernel<D, NS, WITH_GA>, smem))) return rc;                  \
    grad_out_kernel<D, NS, WITH_GA><<<blocks, kGoThreads, smem, st>>>(g, ldg, a, w, N, ga_add, ga, partial)
    switch (C / kGoCols) {
        case 1: B200VQ_GO(1); break;
        case 2: B200VQ_GO(2); break;
        case 3: B200VQ_GO(3); break;
        default: B200VQ_GO(4); break;
    }
#undef B200VQ_GO
    return launch_check();
}
size_t grad_out_smem(int C, int D) {
    switch (D) {
        case 8: return go_smem_bytes<8>(C);
        case 16: return go_smem_bytes<16>(C);
        case 24: return go_smem_bytes<24>(C);
        default: return go_smem_bytes<32>(C);
    }
}
}  // namespace
}  // extern "C++"

namespace {
// partial: nullable workspace of weight_grad_slabs() * (C*D + C) floats -> atomic-free, deterministic
inline int weight_grad_slabs() { return 2 * sm_count(); }
int launch_weight_grad(const float* x, int64_t ldx, const float* y, int64_t N, int C, int D, float* dw, int dw_is_dc,
                       float* colsum, float* partial, cudaStream_t st) {
    if (N < 0 || !b200vq_linear_supported(C, D)) return N < 0 ? B200VQ_EINVAL : B200VQ_EDIM;
    if (!dw) return B200VQ_EINVAL;
    int rc = B200VQ_OK;
    static const bool wg_old = std::getenv("B200VQ_WG_OLD") != nullptr;   // tuning aid: the chunk-outer kernel
    if (partial && N > 0 && !wg_old && (C % kGoCols) == 0 && C <= 4 * kGoCols && x && y && ldx >= C && !(ldx & 3) &&
        !(reinterpret_cast<uintptr_t>(x) & 15) && !(reinterpret_cast<uintptr_t>(y) & 15)) {
        // token-outer kernel: every column's accumulators stay in registers, one slab per SM
        const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((N + kGoTok - 1) / kGoTok, static_cast<int64_t>(sm_count())));
        switch (D) {
            case 8: rc = launch_grad_out_d<8, false>(x, ldx, y, nullptr, N, C, nullptr, nullptr, partial, blocks, st); break;
            case 16: rc = launch_grad_out_d<16, false>(x, ldx, y, nullptr, N, C, nullptr, nullptr, partial, blocks, st); break;
            case 24: rc = launch_grad_out_d<24, false>(x, ldx, y, nullptr, N, C, nullptr, nullptr, partial, blocks, st); break;
            default: rc = launch_grad_out_d<32, false>(x, ldx, y, nullptr, N, C, nullptr, nullptr, partial, blocks, st); break;
        }
        if (rc) return rc;
        const int total = C * D + C;
        weight_grad_reduce_kernel<<<(total + 63) / 64, dim3(64, kWgRedY), 0, st>>>(partial, static_cast<int>(blocks), C, D, dw, dw_is_dc, colsum);
        return launch_check();
    }
    const int64_t tiles = (N + kWgTok - 1) / kWgTok;
    if (tiles < weight_grad_slabs()) partial = nullptr;   // every slab must be written by its CTA
    if (!partial || N == 0) {
        if ((rc = cuda_check(cudaMemsetAsync(dw, 0, static_cast<size_t>(C) * D * sizeof(float), st)))) return rc;
        if (colsum && (rc = cuda_check(cudaMemsetAsync(colsum, 0, static_cast<size_t>(C) * sizeof(float), st)))) return rc;
    }
    if (N == 0) return B200VQ_OK;
    if (!x || !y || ldx < C) return B200VQ_EINVAL;
    if ((ldx & 3) || (reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(y) & 15)) return B200VQ_EALIGN;
    int wg_dbg = 0;
    if (const char* env = std::getenv("B200VQ_WG_DEBUG")) wg_dbg = std::atoi(env);
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>(tiles, static_cast<int64_t>(weight_grad_slabs())));
#define B200VQ_WG(DD)                                                                \
    if ((rc = set_smem(weight_grad_kernel<DD>, wg_smem_bytes<DD>()))) return rc;     \
    weight_grad_kernel<DD><<<blocks, 256, wg_smem_bytes<DD>(), st>>>(x, ldx, y, N, C, dw, dw_is_dc, colsum, partial, wg_dbg)
    switch (D) {
        case 8: B200VQ_WG(8); break;
        case 16: B200VQ_WG(16); break;
        case 24: B200VQ_WG(24); break;
        default: B200VQ_WG(32); break;
    }
#undef B200VQ_WG
    if ((rc = launch_check()) || !partial) return rc;
    const int total = C * D + C;
    weight_grad_reduce_kernel<<<(total + 63) / 64, dim3(64, kWgRedY), 0, st>>>(partial, static_cast<int>(blocks), C, D, dw, dw_is_dc, colsum);
    return launch_check();
}
}  // namespace

int b200vq_weight_grad(const float* x, int64_t ldx, const float* y, int64_t N, int C, int D, float* dw, int dw_is_dc,
                       float* colsum, b200vq_stream_t stream) {
    return launch_weight_grad(x, ldx, y, N, C, D, dw, dw_is_dc, colsum, nullptr, static_cast<cudaStream_t>(stream));
}


int b200vq_grad_out_supported(int C, int D) {
    if (!lin_dim_ok(D) || C < kGoCols || (C % kGoCols) != 0 || C > 4 * kGoCols) return 0;
    return grad_out_smem(C, D) <= 227 * 1024 ? 1 : 0;
}

int b200vq_grad_out(const float* g, int64_t ldg, const float* a, const float* w_out, int64_t N, int C, int D, const float* ga_add,
                    float* ga, float* dw, float* colsum, void* workspace, size_t workspace_bytes, b200vq_stream_t stream) {
    if (N < 0) return B200VQ_EINVAL;
    if (!b200vq_grad_out_supported(C, D)) return B200VQ_EDIM;
    if (!dw || !ga) return B200VQ_EINVAL;
    if (!workspace || workspace_bytes < b200vq_weight_grad_workspace_bytes(C, D)) return B200VQ_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc = B200VQ_OK;
    if (N == 0) {
        if ((rc = cuda_check(cudaMemsetAsync(dw, 0, static_cast<size_t>(C) * D * sizeof(float), st)))) return rc;
        if (colsum && (rc = cuda_check(cudaMemsetAsync(colsum, 0, static_cast<size_t>(C) * sizeof(float), st)))) return rc;
        return B200VQ_OK;
    }
    if (!g || !a || !w_out || ldg < C) return B200VQ_EINVAL;
    if ((ldg & 3) || (reinterpret_cast<uintptr_t>(g) & 15) || (reinterpret_cast<uintptr_t>(a) & 15) ||
        (reinterpret_cast<uintptr_t>(workspace) & 15))
        return B200VQ_EALIGN;
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((N + kGoTok - 1) / kGoTok, static_cast<int64_t>(sm_count())));
    float* partial = static_cast<float*>(workspace);
    switch (D) {
        case 8: rc = launch_grad_out_d<8, true>(g, ldg, a, w_out, N, C, ga_add, ga, partial, blocks, st); break;
        case 16: rc = launch_grad_out_d<16, true>(g, ldg, a, w_out, N, C, ga_add, ga, partial, blocks, st); break;
        case 24: rc = launch_grad_out_d<24, true>(g, ldg, a, w_out, N, C, ga_add, ga, partial, blocks, st); break;
        default: rc = launch_grad_out_d<32, true>(g, ldg, a, w_out, N, C, ga_add, ga, partial, blocks, st); break;
    }
    if (rc) return rc;
    const int total = C * D + C;
    weight_grad_reduce_kernel<<<(total + 63) / 64, dim3(64, kWgRedY), 0, st>>>(partial, static_cast<int>(blocks), C, D, dw, 0, colsum);
    return launch_check();
}

namespace {
__global__ void pack3_kernel(const float* a, const float* b, const float* c, float* out3) {
    if (threadIdx.x == 0) {
        out3[0] = a ? *a : 0.f;
        out3[1] = b ? *b : 0.f;
        out3[2] = c ? *c : 0.f;
    }
}
}  // namespace

int b200vq_pack3(const float* a, const float* b, const float* c, float* out3, b200vq_stream_t stream) {
    if (!out3) return B200VQ_EINVAL;
    pack3_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(a, b, c, out3);
    return launch_check();
}

size_t b200vq_weight_grad_workspace_bytes(int C, int D) {
    if (C <= 0 || D <= 0) return 0;
    return static_cast<size_t>(weight_grad_slabs()) * (static_cast<size_t>(C) * D + C) * sizeof(float);
}

int b200vq_weight_grad_ws(const float* x, int64_t ldx, const float* y, int64_t N, int C, int D, float* dw, int dw_is_dc,
                          float* colsum, void* workspace, size_t workspace_bytes, b200vq_stream_t stream) {
    if (!workspace || workspace_bytes < b200vq_weight_grad_workspace_bytes(C, D)) return B200VQ_EINVAL;
    if (reinterpret_cast<uintptr_t>(workspace) & 15) return B200VQ_EALIGN;
    return launch_weight_grad(x, ldx, y, N, C, D, dw, dw_is_dc, colsum, static_cast<float*>(workspace),
                              static_cast<cudaStream_t>(stream));
}

int b200vq_colsum(const float* y, int64_t N, int D, float* out, b200vq_stream_t stream) {
    if (N < 0 || D <= 0 || D > 32 || !out) return B200VQ_EINVAL;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc = cuda_check(cudaMemsetAsync(out, 0, static_cast<size_t>(D) * sizeof(float), st));
    if (rc || N == 0) return rc;
    if (!y) return B200VQ_EINVAL;
    const int per = 256 / D;
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((N + per - 1) / per, 8ll * sm_count()));
    colsum_small_kernel<<<blocks, 256, 0, st>>>(y, N, D, out);
    return launch_check();
}

}  // extern "C"

// ---- whole-Bottleneck entry points: one host call per direction ------------------------------------
// (reference Bottleneck.forward, models/bottleneck.py:170-188, and its autograd)
namespace {
__global__ void add_inplace_kernel(float* __restrict__ dst, const float* __restrict__ src, int64_t n) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) dst[i] += src[i];
}
inline int add_inplace(float* dst, const float* src, int64_t n, cudaStream_t st) {
    if (n <= 0) return B200VQ_OK;
    add_inplace_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(dst, src, n);
    return launch_check();
}
// mean over the batch of |x[b, 0, :]| and |x[b, n-1, :]| (Bottleneck.forward's diagnostics, bottleneck.py:171-172)
__global__ void __launch_bounds__(128) input_norms_kernel(const float* __restrict__ x, int64_t ldx, int64_t n_tok, int C,
                                                          int64_t B, float* __restrict__ out2) {
    __shared__ float s_part[2][4];
    const float* r0 = x + static_cast<int64_t>(blockIdx.x) * n_tok * ldx;
    const float* r1 = r0 + (n_tok - 1) * ldx;
    float a = 0.f, b = 0.f;
    for (int c = threadIdx.x; c < C; c += 128) {
        const float u = r0[c], v = r1[c];
        a = fmaf(u, u, a);
        b = fmaf(v, v, b);
    }
    a = warp_sum(a);
    b = warp_sum(b);
    if ((threadIdx.x & 31) == 0) { s_part[0][threadIdx.x >> 5] = a; s_part[1][threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x < 2) {
        const float t = s_part[threadIdx.x][0] + s_part[threadIdx.x][1] + s_part[threadIdx.x][2] + s_part[threadIdx.x][3];
        atomicAdd(out2 + threadIdx.x, sqrtf(t) / static_cast<float>(B));
    }
}
struct BnLayout {
    size_t e2, image, vq, ga, gz, gemb, wg, total;
};
inline BnLayout bn_layout(int64_t N, int C, int D, int64_t K) {
    BnLayout l;
    size_t off = 0;
    l.e2 = off; off += align_up(static_cast<size_t>(K) * 4);
    l.image = off; off += align_up(b200vq_tc_image_bytes(K, D));
    l.vq = off; off += align_up(b200vq_workspace_bytes(N, D, K));
    l.ga = off; off += align_up(static_cast<size_t>(N) * D * 4);
    l.gz = off; off += align_up(static_cast<size_t>(N) * D * 4);
    l.gemb = off; off += align_up(static_cast<size_t>(K) * D * 4);
    l.wg = off; off += align_up(static_cast<size_t>(2 * sm_count()) * (static_cast<size_t>(C) * D + C) * 4);
    l.total = off;
    return l;
}
}  // namespace

extern "C" {

size_t b200vq_bottleneck_workspace_bytes(int64_t N, int C, int D, int64_t K) {
    if (N < 0 || K <= 0 || D <= 0 || C <= 0) return 0;
    return bn_layout(N, C, D, K).total;
}

int b200vq_bottleneck_forward(const float* x, int64_t ldx, int64_t N, int C, const float* w_in, const float* b_in,
                              const float* weight, int64_t K, int D, const float* w_out, const float* b_out, int mode,
                              int l2norm, float t_inv, uint64_t seed, uint64_t offset, int64_t n_first, float beta,
                              float w_cb, int use_tc, float* z, float* emb, float* wnorm, float* znorm, float* ste,
                              int64_t* idx, float* zlen, int32_t* hist, int32_t* hist_first, float* losses, float* usage4,
                              float* norms2, float* out, int64_t ldo, void* workspace, size_t workspace_bytes,
                              b200vq_stream_t stream) {
    if (N <= 0 || K <= 0 || !valid_mode(mode)) return B200VQ_EINVAL;
    if (!b200vq_linear_supported(C, D)) return B200VQ_EDIM;
    if (!x || !w_in || !weight || !w_out || !z || !emb || !wnorm || !znorm || !ste || !idx || !zlen || !hist || !hist_first ||
        !losses || !out)
        return B200VQ_EINVAL;
    const BnLayout l = bn_layout(N, C, D, K);
    if (!workspace || workspace_bytes < l.total) return B200VQ_EWORKSPACE;
    if (reinterpret_cast<uintptr_t>(workspace) & 15) return B200VQ_EALIGN;
    char* ws = static_cast<char*>(workspace);
    float* e2 = reinterpret_cast<float*>(ws + l.e2);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc = launch_linear_in(x, ldx, N, C, w_in, 0, b_in, D, z, true, st);
    if (rc) return rc;
    if (norms2) {
        if (n_first <= 0 || N % n_first) return B200VQ_EINVAL;
        if ((rc = cuda_check(cudaMemsetAsync(norms2, 0, 2 * sizeof(float), st)))) return rc;
        input_norms_kernel<<<static_cast<unsigned>(N / n_first), 128, 0, st>>>(x, ldx, n_first, C, N / n_first, norms2);
        if ((rc = launch_check())) return rc;
    }
    if ((rc = b200vq_prepare_codebook(weight, K, D, l2norm, emb, e2, wnorm, stream))) return rc;
    void* image = nullptr;
    if (use_tc && b200vq_tc_supported(D, l2norm, mode)) {
        image = ws + l.image;
        if ((rc = b200vq_prepare_codebook_tc(emb, K, D, image, stream))) return rc;
    }
    rc = b200vq_forward(z, B200VQ_DT_F32, D, N, D, emb, e2, image, K, mode, l2norm, t_inv, seed, offset, n_first, beta, w_cb,
                        znorm, ste, idx, zlen, hist, hist_first, losses, ws + l.vq, b200vq_workspace_bytes(N, D, K), stream);
    if (rc) return rc;
    if (usage4 && (rc = b200vq_usage_stats(hist_first, hist, K, usage4, stream))) return rc;
    return b200vq_linear_out(ste, N, D, w_out, 0, b_out, C, out, ldo, stream);
}

int b200vq_bottleneck_backward(const float* g_out, int64_t ldg, const float* g_losses, const float* g_ste,
                               const float* g_znorm, const float* g_emb, const float* g_z, const float* x, int64_t ldx,
                               int64_t N, int C, const float* w_in, const float* w_out, const float* emb,
                               const float* wnorm, const float* znorm, const float* zlen, const int64_t* idx,
                               const float* ste, int64_t K, int D, int l2norm, float beta, float w_cb, float* grad_x,
                               int64_t ldgx, float* grad_w_in, float* grad_b_in, float* grad_weight, float* grad_w_out,
                               float* grad_b_out, void* workspace, size_t workspace_bytes, b200vq_stream_t stream) {
    if (N <= 0 || K <= 0) return B200VQ_EINVAL;
    if (!b200vq_linear_supported(C, D)) return B200VQ_EDIM;
    if (!x || !w_in || !w_out || !emb || !wnorm || !znorm || !zlen || !idx || !ste) return B200VQ_EINVAL;
    const BnLayout l = bn_layout(N, C, D, K);
    if (!workspace || workspace_bytes < l.total) return B200VQ_EWORKSPACE;
    char* ws = static_cast<char*>(workspace);
    float* ga = reinterpret_cast<float*>(ws + l.ga);      // grad of 'regularized_z'
    float* gz = reinterpret_cast<float*>(ws + l.gz);      // grad of 'projected_z'
    float* gemb = reinterpret_cast<float*>(ws + l.gemb);  // grad of the normalised codebook
    float* wgp = reinterpret_cast<float*>(ws + l.wg);     // per-CTA weight-gradient slabs
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc;
    const float* ga_in = g_ste;
    if (g_out) {
        // grad_a = g_out W_out ; grad_W_out = g_out^T ste ; grad_b_out = column sums of g_out
        static const bool split_env = std::getenv("B200VQ_GRAD_OUT_SPLIT") != nullptr;   // tuning aid: the two-kernel path
        if (grad_w_out && !split_env && b200vq_grad_out_supported(C, D)) {
            // one pass over g_out: grad_a (+ g_ste), grad_W_out and grad_b_out together
            if ((rc = b200vq_grad_out(g_out, ldg, ste, w_out, N, C, D, g_ste, ga, grad_w_out, grad_b_out, wgp,
                                      b200vq_weight_grad_workspace_bytes(C, D), stream)))
                return rc;
        } else {
            if ((rc = launch_linear_in(g_out, ldg, N, C, w_out, 1, nullptr, D, ga, false, st))) return rc;
            if (g_ste && (rc = add_inplace(ga, g_ste, N * D, st))) return rc;
            if (grad_w_out && (rc = launch_weight_grad(g_out, ldg, ste, N, C, D, grad_w_out, 0, grad_b_out, wgp, st))) return rc;
        }
        ga_in = ga;
    } else {
        if (grad_w_out && (rc = cuda_check(cudaMemsetAsync(grad_w_out, 0, static_cast<size_t>(C) * D * 4, st)))) return rc;
        if (grad_b_out && (rc = cuda_check(cudaMemsetAsync(grad_b_out, 0, static_cast<size_t>(C) * 4, st)))) return rc;
    }
    if ((rc = b200vq_backward(ga_in, g_znorm, g_losses, znorm, zlen, idx, emb, nullptr, 0, N, D, K, l2norm, N, beta, w_cb, gz,
                              B200VQ_DT_F32, gemb, stream)))
        return rc;
    if (grad_weight && (rc = b200vq_codebook_backward(gemb, g_emb, emb, wnorm, K, D, l2norm, grad_weight, stream))) return rc;
    if (g_z && (rc = add_inplace(gz, g_z, N * D, st))) return rc;
    if (grad_x && (rc = b200vq_linear_out(gz, N, D, w_in, 1, nullptr, C, grad_x, ldgx, stream))) return rc;
    if (grad_w_in && (rc = launch_weight_grad(x, ldx, gz, N, C, D, grad_w_in, 1, nullptr, wgp, st))) return rc;
    if (grad_b_in && (rc = b200vq_colsum(gz, N, D, grad_b_in, stream))) return rc;
    return B200VQ_OK;
}

}  // extern "C"

// ---- streaming entropy regulariser (vq_entropy.cuh) --------------------------------------------------
namespace {
template <int PASS, int R>
int launch_entropy_r(int D, const EntropyArgs& a, dim3 grid, cudaStream_t st) {
    switch (D) {
        case 8: entropy_pass_kernel<8, PASS, R><<<grid, kEntThreads, 0, st>>>(a); break;
        case 16: entropy_pass_kernel<16, PASS, R><<<grid, kEntThreads, 0, st>>>(a); break;
        case 24: entropy_pass_kernel<24, PASS, R><<<grid, kEntThreads, 0, st>>>(a); break;
        case 32: entropy_pass_kernel<32, PASS, R><<<grid, kEntThreads, 0, st>>>(a); break;
        case 64:   // one owner vector per thread only: two would not fit the register file (entropy_plan)
            if constexpr (R == 1) { entropy_pass_kernel<64, PASS, 1><<<grid, kEntThreads, 0, st>>>(a); break; }
            return B200VQ_EDIM;
        default: return B200VQ_EDIM;
    }
    return launch_check();
}
inline int64_t round_tile(int64_t v) { return (v + kEntTile - 1) / kEntTile * kEntTile; }

// Work split of one pass: owners per thread (2 whenever owners x splits can still fill the machine), CTAs over the
// owners, and splits of the streamed side up to ~8 CTAs per SM (each split keeps >= 4 tiles).
struct EntPlan { int own; dim3 grid; int64_t per; };
inline EntPlan entropy_plan(int64_t n_own, int64_t n_str, int D) {
    EntPlan pl;
    const int64_t sms = sm_count();
    const int64_t max_splits = std::max<int64_t>(1, n_str / (4 * kEntTile));
    const int64_t gx2 = (n_own + 2 * kEntThreads - 1) / (2 * kEntThreads);
    pl.own = (D <= 32 && gx2 * max_splits >= sms) ? 2 : 1;
    const int64_t gx = (n_own + pl.own * kEntThreads - 1) / (pl.own * kEntThreads);
    // ~8 CTAs per SM measured best on B200 (profiles/r01g_summary.md): the passes are issue-latency bound, and a split
    // only costs one atomic per output element (or the ROWSTATS merge)
    static const int target = [] { const char* e = std::getenv("B200VQ_ENT_CTAS_PER_SM"); return e && std::atoi(e) > 0 ? std::atoi(e) : 8; }();
    int64_t splits = std::min<int64_t>(max_splits, std::max<int64_t>(1, (target * sms + gx / 2) / gx));
    splits = std::min<int64_t>(splits, 65535);
    pl.per = round_tile((n_str + splits - 1) / splits);
    splits = (n_str + pl.per - 1) / pl.per;
    pl.grid = dim3(static_cast<unsigned>(gx), static_cast<unsigned>(splits));
    return pl;
}
template <int PASS>
int launch_entropy(int D, EntropyArgs& a, const EntPlan& pl, cudaStream_t st) {
    a.str_per_split = pl.per;
    return pl.own == 2 ? launch_entropy_r<PASS, 2>(D, a, pl.grid, st) : launch_entropy_r<PASS, 1>(D, a, pl.grid, st);
}
}  // namespace

extern "C" {

int b200vq_entropy_supported(int D) { return (D == 8 || D == 16 || D == 24 || D == 32 || D == 64) ? 1 : 0; }

size_t b200vq_entropy_scratch_bytes(int64_t N, int64_t K) {
    if (N <= 0 || K <= 0) return 0;
    // upper bound over D: two owners per thread give the fewest owner CTAs, hence the most splits
    const EntPlan pl = entropy_plan(N, K, 8);
    return align_up(static_cast<size_t>(pl.grid.y) * static_cast<size_t>(N) * 3 * sizeof(float));
}

int b200vq_entropy_forward(const float* z, int64_t N, int D, const float* emb, int64_t K, float t_inv, float* lse,
                           float* ent, float* psum, float* w, double* ent_sum, float* out3, void* scratch,
                           size_t scratch_bytes, b200vq_stream_t stream) {
    if (!b200vq_entropy_supported(D)) return B200VQ_EDIM;
    if (N <= 0 || K <= 0 || !z || !emb || !lse || !ent || !psum || !w || !ent_sum || !out3 || !(t_inv > 0.f))
        return B200VQ_EINVAL;
    if ((reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(emb)) & 15) return B200VQ_EALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc;
    if ((rc = cuda_check(cudaMemsetAsync(psum, 0, static_cast<size_t>(K) * sizeof(float), st)))) return rc;
    if ((rc = cuda_check(cudaMemsetAsync(ent_sum, 0, sizeof(double), st)))) return rc;
    const float inv_n = 1.0f / static_cast<float>(N);
    // rows own, codes stream: lse_i, ent_i, sum of ent
    const EntPlan rows = entropy_plan(N, K, D);
    EntropyArgs a{};
    a.t_inv = t_inv;
    a.inv_n = inv_n;
    a.own = z; a.n_own = N; a.str = emb; a.n_str = K;
    a.out_a = lse; a.out_b = ent; a.out_sum = ent_sum;
    if (rows.grid.y > 1) {
        if (!scratch || scratch_bytes < b200vq_entropy_scratch_bytes(N, K)) return B200VQ_EWORKSPACE;
        a.out_vec = static_cast<float*>(scratch);
    }
    if ((rc = launch_entropy<ENT_ROWSTATS>(D, a, rows, st))) return rc;
    if (rows.grid.y > 1) {
        entropy_merge_kernel<<<static_cast<unsigned>((N + kEntThreads - 1) / kEntThreads), kEntThreads, 0, st>>>(
            static_cast<const float*>(scratch), static_cast<int>(rows.grid.y), N, lse, ent, ent_sum);
        if ((rc = launch_check())) return rc;
    }
    // codes own, rows stream: psum_k
    EntropyArgs b{};
    b.t_inv = t_inv;
    b.inv_n = inv_n;
    b.own = emb; b.n_own = K; b.str = z; b.n_str = N; b.str_a = lse; b.out_a = psum;
    if ((rc = launch_entropy<ENT_COLSUM>(D, b, entropy_plan(K, N, D), st))) return rc;
    entropy_finalize_kernel<<<1, 1024, 0, st>>>(psum, K, inv_n, ent_sum, w, out3);
    return launch_check();
}

int b200vq_entropy_backward(const float* z, int64_t N, int D, const float* emb, int64_t K, float t_inv,
                            const float* lse, const float* ent, const float* w, const float* g3, float* c,
                            float* grad_z, float* grad_emb, b200vq_stream_t stream) {
    if (!b200vq_entropy_supported(D)) return B200VQ_EDIM;
    if (N <= 0 || K <= 0 || !z || !emb || !lse || !ent || !w || !g3 || !c || !(t_inv > 0.f)) return B200VQ_EINVAL;
    if ((reinterpret_cast<uintptr_t>(z) | reinterpret_cast<uintptr_t>(emb) | reinterpret_cast<uintptr_t>(grad_z)) & 15)
        return B200VQ_EALIGN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int rc;
    const float inv_n = 1.0f / static_cast<float>(N);
    const EntPlan rows = entropy_plan(N, K, D);
    EntropyArgs a{};
    a.t_inv = t_inv;
    a.inv_n = inv_n;
    a.own = z; a.n_own = N; a.str = emb; a.n_str = K;
    a.own_a = lse; a.str_a = w; a.out_a = c;
    if (rows.grid.y > 1 && (rc = cuda_check(cudaMemsetAsync(c, 0, static_cast<size_t>(N) * sizeof(float), st)))) return rc;
    if ((rc = launch_entropy<ENT_ROWDOT>(D, a, rows, st))) return rc;
    if (grad_z) {
        if (rows.grid.y > 1 &&
            (rc = cuda_check(cudaMemsetAsync(grad_z, 0, static_cast<size_t>(N) * D * sizeof(float), st))))
            return rc;
        a.own_b = ent; a.own_c = c; a.g3 = g3; a.out_a = nullptr; a.out_vec = grad_z;
        if ((rc = launch_entropy<ENT_GRAD_ROWS>(D, a, rows, st))) return rc;
    }
    if (grad_emb) {
        if ((rc = cuda_check(cudaMemsetAsync(grad_emb, 0, static_cast<size_t>(K) * D * sizeof(float), st)))) return rc;
        EntropyArgs b{};
        b.t_inv = t_inv;
        b.inv_n = inv_n;
        b.own = emb; b.n_own = K; b.str = z; b.n_str = N;
        b.own_a = w; b.str_a = lse; b.str_b = ent; b.str_c = c; b.g3 = g3; b.out_vec = grad_emb;
        if ((rc = launch_entropy<ENT_GRAD_CODES>(D, b, entropy_plan(K, N, D), st))) return rc;
    }
    return B200VQ_OK;
}

}  // extern "C"

// b200vq -- C ABI (include/b200vq.h) over the sm_100a kernels.  No torch types here.
// This unit: library state, the exact-fp32 scoring / finish / backward entry points and the whole-Bottleneck calls.
#include "vq_generic.cuh"
#include "b200vq_host.h"

using namespace b200vq;
using namespace b200vq::host;

namespace b200vq {
namespace host {
thread_local int g_last_cuda_error = 0;
std::atomic<unsigned long long> g_launches{0};
}  // namespace host
}  // namespace b200vq

extern "C" {

int b200vq_version(void) { return B200VQ_VERSION; }
int b200vq_last_cuda_error(void) { return g_last_cuda_error; }
int b200vq_max_dim(void) { return kMaxDim; }
uint64_t b200vq_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

}  // extern "C"

namespace {
__global__ void __launch_bounds__(256) fma_probe_kernel(int64_t iters, float* __restrict__ out) {
    float a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = 1.0f + 1e-3f * static_cast<float>(threadIdx.x + i);
    const float m = 0.999999f, c = 1e-7f * static_cast<float>(blockIdx.x);
    for (int64_t it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], m, c);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i];
    if (s == 123.456f) *out = s;   // keeps the chains live; never true in practice
}
}  // namespace

extern "C" {

int b200vq_fma_probe(int64_t iters, float* out, double* out_flops_per_launch, b200vq_stream_t stream) {
    if (iters <= 0 || !out) return B200VQ_EINVAL;
    const unsigned blocks = static_cast<unsigned>(sm_count()) * 8u;
    fma_probe_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(iters, out);
    if (out_flops_per_launch) *out_flops_per_launch = 2.0 * 8.0 * static_cast<double>(iters) * blocks * 256.0;
    return launch_check();
}

const char* b200vq_error_string(int code) {
    switch (code) {
        case B200VQ_OK: return "ok";
        case B200VQ_EINVAL: return "invalid argument (shape, size or enum)";
        case B200VQ_EDIM: return "unsupported code dimension";
        case B200VQ_EALIGN: return "pointer or offset not aligned as required";
        case B200VQ_ECUDA: return "CUDA error (see b200vq_last_cuda_error)";
        case B200VQ_EWORKSPACE: return "workspace null or too small";
        case B200VQ_ENOALLOC_SIZE: return "size exceeds the library's built-in scratch (B200VQ_NOALLOC_SCRATCH_BYTES); use the workspace variant";
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

}  // extern "C"

namespace {
// b200vq_forward with a device-resident Philox state (nullable `philox_dev`), as the whole-Bottleneck call passes it
int forward_impl(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D, const float* emb, const float* e2,
                 const void* image, int64_t K, int mode, int l2norm, float t_inv, uint64_t seed, uint64_t offset,
                 unsigned long long* philox_dev, int64_t n_first, float beta, float w_cb, float* znorm_out, float* ste_out,
                 int64_t* idx_out, float* zlen_out, int32_t* hist, int32_t* hist_first, float* losses, void* workspace,
                 size_t workspace_bytes, cudaStream_t st) {
    if (N <= 0 || K <= 0) return B200VQ_EINVAL;
    if (!workspace || workspace_bytes < b200vq_workspace_bytes(N, D, K)) return B200VQ_EWORKSPACE;
    if (reinterpret_cast<uintptr_t>(workspace) & 15) return B200VQ_EALIGN;
    uint64_t* keys = static_cast<uint64_t*>(workspace);
    double* scratch = reinterpret_cast<double*>(static_cast<char*>(workspace) + align_up(static_cast<size_t>(N) * 8));
    const bool tc = image && b200vq_tc_supported(D, l2norm, mode);
    int rc;
    // one fill per histogram, always: two buffers that happen to be adjacent in the address space may still be two
    // allocations, and a memset that spans two allocations is an invalid argument (seen with the torch allocator)
    if (hist && (rc = cuda_check(cudaMemsetAsync(hist, 0, static_cast<size_t>(K) * 4, st)))) return rc;
    if (hist_first && (rc = cuda_check(cudaMemsetAsync(hist_first, 0, static_cast<size_t>(K) * 4, st)))) return rc;
    if (tc && mode == B200VQ_MODE_SAMPLE) {
        // keys + sums are zeroed; the deterministic modes' scratch is not needed
        if ((rc = cuda_check(cudaMemsetAsync(workspace, 0, align_up(static_cast<size_t>(N) * 8) + align_up(2 * sizeof(double)), st))))
            return rc;
        rc = host::launch_tc_sample(z, z_dtype, z_row_stride, N, D, image, emb, K, 0, l2norm, t_inv, seed, offset, philox_dev, 0,
                                    keys, nullptr, st);
        if (rc) return rc;
        return b200vq_finish(z, z_dtype, z_row_stride, N, D, emb, K, keys, nullptr, l2norm, n_first, znorm_out, ste_out, idx_out,
                             zlen_out, hist, hist_first, scratch, 1, N, beta, w_cb, losses, st);
    }
    if (philox_dev && mode == B200VQ_MODE_SAMPLE) return B200VQ_EINVAL;   // only the tensor-core sampler reads a device-side state
    if ((rc = cuda_check(cudaMemsetAsync(workspace, 0, b200vq_workspace_bytes(N, D, K), st)))) return rc;
    if (tc) {
        char* base = static_cast<char*>(workspace) + align_up(static_cast<size_t>(N) * 8) + align_up(2 * sizeof(double));
        unsigned int* n_resolved = reinterpret_cast<unsigned int*>(base);
        float* second = reinterpret_cast<float*>(base + align_up(sizeof(unsigned int)));
        rc = b200vq_scores_tc(z, z_dtype, z_row_stride, N, D, image, emb, e2, K, 0, mode, l2norm, t_inv, seed, offset, 0,
                              keys, second, n_resolved, st);
    } else {
        rc = b200vq_scores_partial(z, z_dtype, z_row_stride, N, D, emb, e2, K, 0, mode, l2norm, t_inv, seed, offset, 0,
                                   keys, st);
    }
    if (rc) return rc;
    return b200vq_finish(z, z_dtype, z_row_stride, N, D, emb, K, keys, nullptr, l2norm, n_first, znorm_out, ste_out, idx_out,
                         zlen_out, hist, hist_first, scratch, 1, N, beta, w_cb, losses, st);
}
}  // namespace

extern "C" {

int b200vq_forward(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D, const float* emb,
                   const float* e2, const void* image, int64_t K, int mode, int l2norm, float t_inv, uint64_t seed, uint64_t offset,
                   int64_t n_first, float beta, float w_cb, float* znorm_out, float* ste_out, int64_t* idx_out,
                   float* zlen_out, int32_t* hist, int32_t* hist_first, float* losses, void* workspace,
                   size_t workspace_bytes, b200vq_stream_t stream) {
    return forward_impl(z, z_dtype, z_row_stride, N, D, emb, e2, image, K, mode, l2norm, t_inv, seed, offset, nullptr, n_first,
                        beta, w_cb, znorm_out, ste_out, idx_out, zlen_out, hist, hist_first, losses, workspace, workspace_bytes,
                        static_cast<cudaStream_t>(stream));
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
// ---- side lane: independent small kernels of the one-call step run beside the big streaming passes ------------
// One non-blocking stream + a few events per device, created on the first call that is not under stream capture and
// kept for the life of the process.  fork() makes the lane wait for everything enqueued on the caller's stream so
// far, join() makes the caller's stream wait for the lane; both are plain event record / wait, so the calls stay
// asynchronous and capturable (the lane joins a CUDA-graph capture through the event and leaves it at join()).
// B200VQ_NO_OVERLAP=1 keeps everything on the caller's stream.
struct SideLane {
    std::atomic<int> state{0};   // 0 = not tried, 1 = ready, -1 = unavailable
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
};
SideLane g_side[64];
std::mutex g_side_mu;

// Which forks are used: bit 1 codebook preparation beside in_linear, 2 usage statistics beside out_linear, 4 backward
// tail (codebook Jacobian, bias gradient, grad_W_in) beside grad_x, 8 grad_W_in stays on the caller's stream.
// The streaming kernels run one persistent CTA per SM, so a side kernel that takes an SM stretches them when both
// really start together, which is what a replayed graph does: under capture only the forks that pay there are used
// (measured, profiles/r02_summary.md: 13 -> 0.554 ms against 0.569-0.576 serial; eager 7 -> 0.570 against 0.604).
// B200VQ_OVERLAP overrides both.
// Outside a capture every fork costs a few host microseconds of event calls: small batches (host-bound) stay serial.
inline int overlap_mask(cudaStream_t st, int64_t N) {
    static const int forced = [] { const char* e = std::getenv("B200VQ_OVERLAP"); return e ? std::atoi(e) : -1; }();
    if (forced >= 0) return forced;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) { cudaGetLastError(); return 0; }
    if (cs != cudaStreamCaptureStatusNone) return N >= 32768 ? 13 : 7;   // small grids leave SMs free: every fork pays
    return N >= 32768 ? 7 : 0;
}
SideLane* side_lane(cudaStream_t st) {
    static const bool off = std::getenv("B200VQ_NO_OVERLAP") != nullptr;
    const int dev = current_device();
    if (off || dev >= 64) return nullptr;
    SideLane& s = g_side[dev];
    const int state = s.state.load(std::memory_order_acquire);
    if (state == 1) return &s;
    if (state < 0) return nullptr;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (cs != cudaStreamCaptureStatusNone) return nullptr;   // never create objects inside a capture: this call runs serially
    std::lock_guard<std::mutex> lock(g_side_mu);
    if (s.state.load(std::memory_order_acquire) == 1) return &s;
    bool ok = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; ok && i < 4; ++i) ok = cudaEventCreateWithFlags(&s.ev[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) cudaGetLastError();
    s.state.store(ok ? 1 : -1, std::memory_order_release);
    return ok ? &s : nullptr;
}
// every early return between fork() and join() still joins (destructor), so a capture is never left forked
struct SideScope {
    SideLane* lane;
    cudaStream_t main;
    int ev_join;
    bool forked = false;
    SideScope(SideLane* l, cudaStream_t m, int ev_fork, int ev_join_) : lane(l), main(m), ev_join(ev_join_) {
        if (!lane) return;
        std::lock_guard<std::mutex> lock(g_side_mu);   // record + wait as one step: another thread may use the same lane
        if (cudaEventRecord(lane->ev[ev_fork], main) == cudaSuccess && cudaStreamWaitEvent(lane->stream, lane->ev[ev_fork], 0) == cudaSuccess)
            forked = true;
        else { cudaGetLastError(); lane = nullptr; }
    }
    cudaStream_t side() const { return forked ? lane->stream : main; }
    int join() {
        if (!forked) return B200VQ_OK;
        forked = false;
        std::lock_guard<std::mutex> lock(g_side_mu);
        int rc = cuda_check(cudaEventRecord(lane->ev[ev_join], lane->stream));
        if (!rc) rc = cuda_check(cudaStreamWaitEvent(main, lane->ev[ev_join], 0));
        return rc;
    }
    ~SideScope() { join(); }
};

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
                              float w_cb, int use_tc, void* z, float* emb, float* wnorm, float* znorm, float* ste,
                              int64_t* idx, float* zlen, int32_t* hist, int32_t* hist_first, float* losses, float* usage4,
                              float* norms2, void* out, int64_t ldo, void* workspace, size_t workspace_bytes,
                              uint64_t* philox_dev, int act_dtype, const float* add, int64_t add_rows,
                              b200vq_stream_t stream) {
    if (N <= 0 || K <= 0 || !valid_mode(mode) || !valid_dtype(act_dtype)) return B200VQ_EINVAL;
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
    if (norms2 && (n_first <= 0 || N % n_first)) return B200VQ_EINVAL;
    SideLane* lane = side_lane(st);
    const int forks = overlap_mask(st, N);
    int rc = B200VQ_OK;
    void* image = nullptr;
    {
        // codebook preparation and the input-norm diagnostics do not depend on the latents: beside in_linear
        SideScope sc((forks & 1) ? lane : nullptr, st, 0, 1);
        cudaStream_t s2 = sc.side();
        if (norms2) {
            if ((rc = cuda_check(cudaMemsetAsync(norms2, 0, 2 * sizeof(float), s2)))) return rc;
            input_norms_kernel<<<static_cast<unsigned>(N / n_first), 128, 0, s2>>>(x, ldx, n_first, C, N / n_first, norms2);
            if ((rc = launch_check())) return rc;
        }
        if ((rc = b200vq_prepare_codebook(weight, K, D, l2norm, emb, e2, wnorm, s2))) return rc;
        if (use_tc && b200vq_tc_supported(D, l2norm, mode)) {
            image = ws + l.image;
            if ((rc = b200vq_prepare_codebook_tc(emb, K, D, image, s2))) return rc;
        }
        if ((rc = launch_linear_in(x, ldx, N, C, w_in, 0, b_in, D, z, act_dtype, true, st))) return rc;
        if ((rc = sc.join())) return rc;
    }
    rc = forward_impl(z, act_dtype, D, N, D, emb, e2, image, K, mode, l2norm, t_inv, seed, offset,
                      reinterpret_cast<unsigned long long*>(philox_dev), n_first, beta, w_cb, znorm, ste, idx, zlen, hist,
                      hist_first, losses, ws + l.vq, b200vq_workspace_bytes(N, D, K), st);
    if (rc) return rc;
    // the usage statistics (one CTA, latency-bound) beside out_linear
    SideScope sc(usage4 && (forks & 2) ? lane : nullptr, st, 2, 3);
    if (usage4 && (rc = b200vq_usage_stats(hist_first, hist, K, usage4, sc.side()))) return rc;
    if ((rc = launch_linear_out(ste, N, D, w_out, 0, b_out, C, out, act_dtype, ldo, nullptr, 0, 0, add, add_rows, nullptr, st))) return rc;
    return sc.join();
}

int b200vq_bottleneck_backward(const void* g_out, int act_dtype, int64_t ldg, const float* g_losses, const float* g_ste,
                               const float* g_znorm, const float* g_emb, const float* g_z, const float* x, int64_t ldx,
                               int64_t N, int C, const float* w_in, const float* w_out, const float* emb,
                               const float* wnorm, const float* znorm, const float* zlen, const int64_t* idx,
                               const float* ste, int64_t K, int D, int l2norm, float beta, float w_cb, float* grad_x,
                               int64_t ldgx, float* grad_w_in, float* grad_b_in, float* grad_weight, float* grad_w_out,
                               float* grad_b_out, void* workspace, size_t workspace_bytes, void* grads_ready_event,
                               b200vq_stream_t stream) {
    if (N <= 0 || K <= 0 || !valid_dtype(act_dtype)) return B200VQ_EINVAL;
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
            if ((rc = launch_grad_out(g_out, act_dtype, ldg, ste, w_out, N, C, D, g_ste, ga, grad_w_out, grad_b_out, wgp,
                                      b200vq_weight_grad_workspace_bytes(C, D), st)))
                return rc;
        } else {
            if (act_dtype != B200VQ_DT_F32) return B200VQ_EDIM;   // the two-kernel path reads fp32 gradients only
            const float* g32 = static_cast<const float*>(g_out);
            if ((rc = launch_linear_in(g32, ldg, N, C, w_out, 1, nullptr, D, ga, B200VQ_DT_F32, false, st))) return rc;
            if (g_ste && (rc = add_inplace(ga, g_ste, N * D, st))) return rc;
            if (grad_w_out && (rc = launch_weight_grad(g32, ldg, ste, N, C, D, grad_w_out, 0, grad_b_out, wgp, st))) return rc;
        }
        ga_in = ga;
    } else {
        if (grad_w_out && (rc = cuda_check(cudaMemsetAsync(grad_w_out, 0, static_cast<size_t>(C) * D * 4, st)))) return rc;
        if (grad_b_out && (rc = cuda_check(cudaMemsetAsync(grad_b_out, 0, static_cast<size_t>(C) * 4, st)))) return rc;
    }
    if ((rc = b200vq_backward(ga_in, g_znorm, g_losses, znorm, zlen, idx, emb, nullptr, 0, N, D, K, l2norm, N, beta, w_cb, gz,
                              B200VQ_DT_F32, gemb, stream)))
        return rc;
    if (g_z && (rc = add_inplace(gz, g_z, N * D, st))) return rc;
    // grad_x (one pass of stores) on the caller's stream; the codebook Jacobian, the bias gradient and grad_W_in (one
    // pass of loads over x) beside it
    const int forks = overlap_mask(st, N);
    SideScope sc(grad_x && (forks & 4) ? side_lane(st) : nullptr, st, 0, 1);
    cudaStream_t s2 = sc.side();
    if (grad_weight && (rc = b200vq_codebook_backward(gemb, g_emb, emb, wnorm, K, D, l2norm, grad_weight, s2))) return rc;
    if (grad_b_in && (rc = b200vq_colsum(gz, N, D, grad_b_in, s2))) return rc;
    // with a ready-event every parameter gradient goes to the lane (or stays on the caller's stream when there is none),
    // so that ONE record covers them all
    cudaStream_t wg_stream = (forks & 8) && !grads_ready_event ? st : s2;
    if (grad_w_in && (rc = launch_weight_grad(x, ldx, gz, N, C, D, grad_w_in, 1, nullptr, wgp, wg_stream))) return rc;
    if (grads_ready_event && (rc = cuda_check(cudaEventRecord(static_cast<cudaEvent_t>(grads_ready_event), s2)))) return rc;
    if (grad_x && (rc = b200vq_linear_out(gz, N, D, w_in, 1, nullptr, C, grad_x, ldgx, stream))) return rc;
    return sc.join();
}

}  // extern "C"

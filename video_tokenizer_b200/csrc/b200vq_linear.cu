// b200vq -- C ABI (include/b200vq.h) over the sm_100a kernels.  No torch types here.
// This unit: the Bottleneck's skinny linears and their gradients.
#include "vq_linear.cuh"
#include "vq_gradout.cuh"
#include "b200vq_host.h"

using namespace b200vq;
using namespace b200vq::host;

extern "C" {

int b200vq_linear_supported(int C, int D) {
    if (!lin_dim_ok(D) || C < 16 || (C % 16) != 0) return 0;
    const size_t in_smem = static_cast<size_t>(D) * (C + 16) * 4, out_smem = static_cast<size_t>(C) * (D + 5) * 4;
    return in_smem <= 200 * 1024 && out_smem <= 200 * 1024 ? 1 : 0;
}

}  // extern "C"
namespace b200vq {
namespace host {
int launch_linear_in(const float* x, int64_t ldx, int64_t N, int C, const float* w, int w_is_cd, const float* bias, int D,
                     void* y, int y_dtype, bool precise, cudaStream_t st) {
    if (N < 0 || !valid_dtype(y_dtype)) return B200VQ_EINVAL;
    if (!b200vq_linear_supported(C, D)) return B200VQ_EDIM;
    if (N == 0) return B200VQ_OK;
    if (!x || !w || !y || ldx < C) return B200VQ_EINVAL;
    if ((ldx & 3) || (reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(w) & 15) ||
        (reinterpret_cast<uintptr_t>(y) & 7))
        return B200VQ_EALIGN;
    const size_t w_smem = static_cast<size_t>(D) * (C + 16) * 4;
    static const bool no_ring = std::getenv("B200VQ_LIN_IN_NO_RING") != nullptr;   // tuning aid
    const bool ring = !no_ring && w_smem + lin_in_ring_bytes() <= 227 * 1024;
    const size_t smem = w_smem + (ring ? lin_in_ring_bytes() : 0);
    // small batches: split the columns over the warps of a CTA until the grid covers most of the SMs
    int ksplit = 1;
    if (ring) {
        static const int forced = [] { const char* e = std::getenv("B200VQ_LIN_IN_KSPLIT"); return e ? std::atoi(e) : 0; }();   // tuning aid
        while (ksplit < 8 && ((C / 16) % (2 * ksplit)) == 0 && (N + kLinInRows / ksplit - 1) / (kLinInRows / ksplit) < 3 * sm_count() / 4) ksplit *= 2;
        if ((forced == 1 || forced == 2 || forced == 4 || forced == 8) && ((C / 16) % forced) == 0) ksplit = forced;
    }
    // (measured: 147 tiles of 448 rows on 65 536 tokens are SLOWER than 128 tiles of 512 -- 56.4 vs 51.9 us -- so the
    // CTA keeps its 16 warps; the kernel takes any whole number of warps)
    const int nthr = kLinInThreads;
    const int64_t rpc = nthr / ksplit;
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((N + rpc - 1) / rpc, static_cast<int64_t>(sm_count())));
    int rc = B200VQ_OK;
#define B200VQ_LIN_IN_T(DD, PP, OT)                                                              \
    {                                                                                            \
        if (ring) {                                                                              \
            if ((rc = set_smem(linear_in_kernel<DD, PP, OT, true>, smem))) return rc;            \
            linear_in_kernel<DD, PP, OT, true><<<blocks, nthr, smem, st>>>(x, ldx, N, C, w, w_is_cd, bias, static_cast<OT*>(y), ksplit); \
        } else {                                                                                 \
            if ((rc = set_smem(linear_in_kernel<DD, PP, OT>, smem))) return rc;                  \
            linear_in_kernel<DD, PP, OT><<<blocks, nthr, smem, st>>>(x, ldx, N, C, w, w_is_cd, bias, static_cast<OT*>(y)); \
        }                                                                                        \
    }
    // fp32 output: 3xTF32 when `precise` (the latents decide the code indices); the 2-byte outputs of an autocast
    // region carry 8 / 11 significant bits, so one TF32 pass (11 bits in, fp32 accumulate) is already finer than them
#define B200VQ_LIN_IN(DD)                                                                        \
    if (y_dtype == B200VQ_DT_F32) {                                                              \
        if (precise) B200VQ_LIN_IN_T(DD, true, float) else B200VQ_LIN_IN_T(DD, false, float)     \
    } else if (y_dtype == B200VQ_DT_BF16) B200VQ_LIN_IN_T(DD, false, __nv_bfloat16)              \
    else B200VQ_LIN_IN_T(DD, false, __half)
    switch (D) {
        case 8: B200VQ_LIN_IN(8); break;
        case 16: B200VQ_LIN_IN(16); break;
        case 24: B200VQ_LIN_IN(24); break;
        default: B200VQ_LIN_IN(32); break;
    }
#undef B200VQ_LIN_IN_T
#undef B200VQ_LIN_IN
    return launch_check();
}
}  // namespace host
}  // namespace b200vq
extern "C" {

int b200vq_linear_in(const float* x, int64_t ldx, int64_t N, int C, const float* w, const float* bias, int D, float* y,
                     b200vq_stream_t stream) {
    return launch_linear_in(x, ldx, N, C, w, 0, bias, D, y, B200VQ_DT_F32, true, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
namespace b200vq {
namespace host {
int launch_linear_out(const float* a, int64_t N, int D, const float* w, int w_is_dc, const float* bias, int C, void* y,
                      int y_dtype, int64_t ldy, const int64_t* gidx, int64_t gK, int gl2, const float* add, int64_t add_rows,
                      int* bad_flag, cudaStream_t st) {
    if (N < 0 || !valid_dtype(y_dtype)) return B200VQ_EINVAL;
    if (!b200vq_linear_supported(C, D)) return B200VQ_EDIM;
    if (N == 0) return B200VQ_OK;
    if (!a || !w || !y || ldy < C) return B200VQ_EINVAL;
    // rows are written 4 elements at a time: 16-byte (fp32) / 8-byte (bf16, fp16) aligned
    if ((ldy & 3) || (reinterpret_cast<uintptr_t>(y) & (y_dtype == B200VQ_DT_F32 ? 15 : 7)) || (reinterpret_cast<uintptr_t>(w) & 15))
        return B200VQ_EALIGN;
    if (add && (add_rows <= 0 || (reinterpret_cast<uintptr_t>(add) & 15))) return add_rows <= 0 ? B200VQ_EINVAL : B200VQ_EALIGN;
    // column-stationary kernel (weights in registers, no shared memory) whenever 16 warps x 4 groups cover the columns
    static const bool lo_old = std::getenv("B200VQ_LO_OLD") != nullptr;   // tuning aid: the shared-memory kernel
    const int ngroups = C / 16, G = (ngroups + kLinOutThreads / 32 - 1) / (kLinOutThreads / 32);
    if (!gidx && !lo_old && G >= 1 && G <= 4 && !(reinterpret_cast<uintptr_t>(a) & 7)) {
        const unsigned cblocks = static_cast<unsigned>(std::min<int64_t>((N + 15) / 16, static_cast<int64_t>(sm_count())));
        int rc2 = B200VQ_OK;
#define B200VQ_LO_COLS(DD, GG, OT) \
        linear_out_cols_kernel<DD, GG, OT><<<cblocks, kLinOutThreads, 0, st>>>(a, N, w, w_is_dc, bias, C, static_cast<OT*>(y), ldy, add, add_rows)
#define B200VQ_LO_COLS_G(DD, OT)                                  \
        switch (G) {                                              \
            case 1: B200VQ_LO_COLS(DD, 1, OT); break;             \
            case 2: B200VQ_LO_COLS(DD, 2, OT); break;             \
            case 3: B200VQ_LO_COLS(DD, 3, OT); break;             \
            default: B200VQ_LO_COLS(DD, 4, OT); break;            \
        }
#define B200VQ_LO_COLS_T(DD)                                                       \
        if (y_dtype == B200VQ_DT_F32) { B200VQ_LO_COLS_G(DD, float) }              \
        else if (y_dtype == B200VQ_DT_BF16) { B200VQ_LO_COLS_G(DD, __nv_bfloat16) } \
        else { B200VQ_LO_COLS_G(DD, __half) }
        switch (D) {
            case 8: B200VQ_LO_COLS_T(8); break;
            case 16: B200VQ_LO_COLS_T(16); break;
            case 24: B200VQ_LO_COLS_T(24); break;
            default: B200VQ_LO_COLS_T(32); break;
        }
#undef B200VQ_LO_COLS_T
#undef B200VQ_LO_COLS_G
#undef B200VQ_LO_COLS
        (void)rc2;
        return launch_check();
    }
    const size_t smem = static_cast<size_t>(C) * (D + 5) * 4;
    int lo_dbg = 0, lo_grid = 1;   // one 16-warp CTA per SM: the weight staging is paid once per SM
    if (const char* env = std::getenv("B200VQ_LO_DEBUG")) lo_dbg = std::atoi(env);
    if (const char* env = std::getenv("B200VQ_LO_GRID")) lo_grid = std::max(1, std::atoi(env));
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((N + kLinRows - 1) / kLinRows, static_cast<int64_t>(lo_grid) * sm_count()));
    int rc = B200VQ_OK;
#define B200VQ_LIN_OUT_T(DD, OT)                                                                  \
    {                                                                                             \
        if ((rc = set_smem(linear_out_kernel<DD, OT>, smem))) return rc;                          \
        linear_out_kernel<DD, OT><<<blocks, kLinOutThreads, smem, st>>>(a, N, w, w_is_dc, bias, C, static_cast<OT*>(y), ldy, lo_dbg, \
                                                                        gidx, gK, gl2, add, add_rows, bad_flag);                  \
    }
#define B200VQ_LIN_OUT(DD)                                                           \
    if (y_dtype == B200VQ_DT_F32) B200VQ_LIN_OUT_T(DD, float)                        \
    else if (y_dtype == B200VQ_DT_BF16) B200VQ_LIN_OUT_T(DD, __nv_bfloat16)          \
    else B200VQ_LIN_OUT_T(DD, __half)
    switch (D) {
        case 8: B200VQ_LIN_OUT(8); break;
        case 16: B200VQ_LIN_OUT(16); break;
        case 24: B200VQ_LIN_OUT(24); break;
        default: B200VQ_LIN_OUT(32); break;
    }
#undef B200VQ_LIN_OUT
#undef B200VQ_LIN_OUT_T
    return launch_check();
}
}  // namespace host
}  // namespace b200vq
extern "C" {

int b200vq_linear_out(const float* a, int64_t N, int D, const float* w, int w_is_dc, const float* bias, int C, float* y,
                      int64_t ldy, b200vq_stream_t stream) {
    return launch_linear_out(a, N, D, w, w_is_dc, bias, C, y, B200VQ_DT_F32, ldy, nullptr, 0, 0, nullptr, 0, nullptr,
                             static_cast<cudaStream_t>(stream));
}

int b200vq_decode_project(const int64_t* idx, int64_t n, const float* weight, int64_t K, int D, int l2norm, const float* w_out,
                          const float* b_out, int C, const float* add, int64_t add_rows, float* out, int64_t ldo, int* bad_flag,
                          b200vq_stream_t stream) {
    if (n < 0 || K <= 0) return B200VQ_EINVAL;
    if (n > 0 && !idx) return B200VQ_EINVAL;
    return launch_linear_out(weight, n, D, w_out, 0, b_out, C, out, B200VQ_DT_F32, ldo, idx, K, l2norm, add, add_rows, bad_flag,
                             static_cast<cudaStream_t>(stream));
}

extern "C++" {
namespace {
template <int D, bool WITH_GA, typename GT = float>
int launch_grad_out_d(const GT* g, int64_t ldg, const float* a, const float* w, int64_t N, int C, const float* ga_add, float* ga,
                      float* partial, unsigned blocks, cudaStream_t st) {
    const size_t smem = WITH_GA ? go_smem_bytes<D>(C) : go_smem_bytes_wgrad<D>();
    int rc = B200VQ_OK;
#define B200VQ_GO(NS)                                                                       \
    if ((rc = set_smem(grad_out_kernel<D, NS, WITH_GA, GT>, smem))) return rc;              \
    grad_out_kernel<D, NS, WITH_GA, GT><<<blocks, kGoThreads, smem, st>>>(g, ldg, a, w, N, ga_add, ga, partial)
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

}  // extern "C"
namespace {
// partial: nullable workspace of weight_grad_slabs() * (C*D + C) floats -> atomic-free, deterministic
inline int weight_grad_slabs() { return 2 * sm_count(); }
}  // namespace
namespace b200vq {
namespace host {
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
}  // namespace host
}  // namespace b200vq
extern "C" {

int b200vq_weight_grad(const float* x, int64_t ldx, const float* y, int64_t N, int C, int D, float* dw, int dw_is_dc,
                       float* colsum, b200vq_stream_t stream) {
    return launch_weight_grad(x, ldx, y, N, C, D, dw, dw_is_dc, colsum, nullptr, static_cast<cudaStream_t>(stream));
}


int b200vq_grad_out_supported(int C, int D) {
    if (!lin_dim_ok(D) || C < kGoCols || (C % kGoCols) != 0 || C > 4 * kGoCols) return 0;
    return grad_out_smem(C, D) <= 227 * 1024 ? 1 : 0;
}

}  // extern "C"
namespace b200vq {
namespace host {
int launch_grad_out(const void* g, int g_dtype, int64_t ldg, const float* a, const float* w_out, int64_t N, int C, int D,
                    const float* ga_add, float* ga, float* dw, float* colsum, void* workspace, size_t workspace_bytes,
                    cudaStream_t st) {
    if (N < 0 || !valid_dtype(g_dtype)) return B200VQ_EINVAL;
    if (!b200vq_grad_out_supported(C, D)) return B200VQ_EDIM;
    if (!dw || !ga) return B200VQ_EINVAL;
    if (!workspace || workspace_bytes < b200vq_weight_grad_workspace_bytes(C, D)) return B200VQ_EINVAL;
    int rc = B200VQ_OK;
    if (N == 0) {
        if ((rc = cuda_check(cudaMemsetAsync(dw, 0, static_cast<size_t>(C) * D * sizeof(float), st)))) return rc;
        if (colsum && (rc = cuda_check(cudaMemsetAsync(colsum, 0, static_cast<size_t>(C) * sizeof(float), st)))) return rc;
        return B200VQ_OK;
    }
    if (!g || !a || !w_out || ldg < C) return B200VQ_EINVAL;
    // rows of g are fetched 16 bytes at a time
    if ((ldg & (g_dtype == B200VQ_DT_F32 ? 3 : 7)) || (reinterpret_cast<uintptr_t>(g) & 15) || (reinterpret_cast<uintptr_t>(a) & 15) ||
        (reinterpret_cast<uintptr_t>(workspace) & 15))
        return B200VQ_EALIGN;
    const unsigned blocks = static_cast<unsigned>(std::min<int64_t>((N + kGoTok - 1) / kGoTok, static_cast<int64_t>(sm_count())));
    float* partial = static_cast<float*>(workspace);
#define B200VQ_GO_T(GT)                                                                                                       \
    switch (D) {                                                                                                              \
        case 8: rc = launch_grad_out_d<8, true, GT>(static_cast<const GT*>(g), ldg, a, w_out, N, C, ga_add, ga, partial, blocks, st); break;   \
        case 16: rc = launch_grad_out_d<16, true, GT>(static_cast<const GT*>(g), ldg, a, w_out, N, C, ga_add, ga, partial, blocks, st); break; \
        case 24: rc = launch_grad_out_d<24, true, GT>(static_cast<const GT*>(g), ldg, a, w_out, N, C, ga_add, ga, partial, blocks, st); break; \
        default: rc = launch_grad_out_d<32, true, GT>(static_cast<const GT*>(g), ldg, a, w_out, N, C, ga_add, ga, partial, blocks, st); break; \
    }
    if (g_dtype == B200VQ_DT_F32) { B200VQ_GO_T(float) }
    else if (g_dtype == B200VQ_DT_BF16) { B200VQ_GO_T(__nv_bfloat16) }
    else { B200VQ_GO_T(__half) }
#undef B200VQ_GO_T
    if (rc) return rc;
    const int total = C * D + C;
    weight_grad_reduce_kernel<<<(total + 63) / 64, dim3(64, kWgRedY), 0, st>>>(partial, static_cast<int>(blocks), C, D, dw, 0, colsum);
    return launch_check();
}
}  // namespace host
}  // namespace b200vq
extern "C" {

int b200vq_grad_out(const float* g, int64_t ldg, const float* a, const float* w_out, int64_t N, int C, int D, const float* ga_add,
                    float* ga, float* dw, float* colsum, void* workspace, size_t workspace_bytes, b200vq_stream_t stream) {
    return launch_grad_out(g, B200VQ_DT_F32, ldg, a, w_out, N, C, D, ga_add, ga, dw, colsum, workspace, workspace_bytes,
                           static_cast<cudaStream_t>(stream));
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

}  // extern "C"

// ---- allocation-free variant of b200vq_weight_grad_ws ---------------------------------------------------------
// The per-CTA slabs live in a file-scope device array instead of a caller workspace: nothing is allocated by the
// call, so it can be used where cudaMalloc is forbidden (stream capture, serving loops).  Every GPU that loads the
// library holds B200VQ_NOALLOC_SCRATCH_BYTES until the process exits.  In host code `g_noalloc_scratch` is only a
// handle of the symbol, never a device address: its address on the current device comes from cudaGetSymbolAddress,
// looked up once per device and cached.
namespace {
__device__ __align__(256) float g_noalloc_scratch[B200VQ_NOALLOC_SCRATCH_BYTES / sizeof(float)];

int noalloc_scratch(float** out) {
    static std::atomic<float*> cached[64];
    const int dev = current_device();   // the lookup is made with the caller's device current
    float* p = dev < 64 ? cached[dev].load(std::memory_order_acquire) : nullptr;
    if (!p) {
        void* sym = nullptr;
        if (int rc = cuda_check(cudaGetSymbolAddress(&sym, g_noalloc_scratch))) return rc;
        if (!sym) return B200VQ_ECUDA;
        p = static_cast<float*>(sym);
        if (dev < 64) cached[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return B200VQ_OK;
}
}  // namespace

extern "C" {

int b200vq_weight_grad_noalloc(const float* x, int64_t ldx, const float* y, int64_t N, int C, int D, float* dw,
                               int dw_is_dc, float* colsum, b200vq_stream_t stream) {
    if (N < 0) return B200VQ_EINVAL;
    if (!b200vq_linear_supported(C, D)) return B200VQ_EDIM;
    // same slab layout as b200vq_weight_grad_ws; (C, D) whose slabs do not fit the built-in array are refused
    if (b200vq_weight_grad_workspace_bytes(C, D) > static_cast<size_t>(B200VQ_NOALLOC_SCRATCH_BYTES)) return B200VQ_ENOALLOC_SIZE;
    float* scratch = nullptr;
    if (int rc = noalloc_scratch(&scratch)) return rc;
    // nothing to clear: every slab that the reduction reads is written in full by its CTA in the same call, and the
    // paths that accumulate with atomics instead zero dw / colsum (caller memory) themselves
    return launch_weight_grad(x, ldx, y, N, C, D, dw, dw_is_dc, colsum, scratch, static_cast<cudaStream_t>(stream));
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

// b200vq -- host-side helpers shared by the translation units of libb200vq.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "b200vq.h"

namespace b200vq {
namespace host {

extern thread_local int g_last_cuda_error;            // b200vq.cu
extern std::atomic<unsigned long long> g_launches;    // b200vq.cu

// B200VQ_TRACE=1: every failing CUDA call / launch is reported on stderr with its source location
inline bool trace_enabled() {
    static const bool on = std::getenv("B200VQ_TRACE") != nullptr;
    return on;
}
inline int cuda_check_at(cudaError_t e, const char* file, int line) {
    if (e == cudaSuccess) return B200VQ_OK;
    g_last_cuda_error = static_cast<int>(e);
    if (trace_enabled()) std::fprintf(stderr, "[b200vq] %s:%d: %s (%d)\n", file, line, cudaGetErrorString(e), static_cast<int>(e));
    cudaGetLastError();  // clear the sticky launch error so later calls can proceed
    return B200VQ_ECUDA;
}
inline int launch_check_at(const char* file, int line) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return cuda_check_at(cudaPeekAtLastError(), file, line);
}
#define cuda_check(e) cuda_check_at((e), __FILE__, __LINE__)
#define launch_check() launch_check_at(__FILE__, __LINE__)

inline int current_device() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) return 0;
    return dev;
}

inline int sm_count() {
    static int cached[64] = {0};
    const int dev = current_device();
    if (dev >= 64) return 148;
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
inline bool lin_dim_ok(int D) { return D == 8 || D == 16 || D == 24 || D == 32; }

constexpr size_t kAlign = 256;
inline size_t align_up(size_t v) { return (v + kAlign - 1) / kAlign * kAlign; }

// dynamic shared memory above 48 KB is opted into per kernel AND per device: set on every launch
template <typename K>
inline int set_smem(K kernel, size_t bytes) {
    if (bytes <= 48 * 1024) return B200VQ_OK;
    return cuda_check(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)));
}

// ---- entry points of one translation unit used by another (b200vq_linear.cu -> b200vq.cu) ----------
// y_dtype / g_dtype: B200VQ_DT_* of the activation the caller's autocast region uses (fp32 outside autocast)
int launch_linear_in(const float* x, int64_t ldx, int64_t N, int C, const float* w, int w_is_cd, const float* bias, int D,
                     void* y, int y_dtype, bool precise, cudaStream_t st);
int launch_linear_out(const float* a, int64_t N, int D, const float* w, int w_is_dc, const float* bias, int C, void* y,
                      int y_dtype, int64_t ldy, const int64_t* gidx, int64_t gK, int gl2, const float* add, int64_t add_rows,
                      int* bad_flag, cudaStream_t st);
int launch_grad_out(const void* g, int g_dtype, int64_t ldg, const float* a, const float* w_out, int64_t N, int C, int D,
                    const float* ga_add, float* ga, float* dw, float* colsum, void* workspace, size_t workspace_bytes,
                    cudaStream_t st);
int launch_weight_grad(const float* x, int64_t ldx, const float* y, int64_t N, int C, int D, float* dw, int dw_is_dc,
                       float* colsum, float* partial, cudaStream_t st);

// ---- b200vq_tc.cu -> b200vq.cu: the tensor-core sampler ---------------------------------------------------------
// philox_dev: nullable device {seed, offset}; when given it overrides seed / offset and the offset is advanced by 4 on
// the device (a one-thread kernel behind the sampler), the GPU-side twin of functional.philox_state.
int launch_tc_sample(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D, const void* image, const float* emb,
                     int64_t K, int64_t k_offset, int l2norm, float t_inv, uint64_t seed, uint64_t offset,
                     unsigned long long* philox_dev, int64_t row_offset, uint64_t* keys, unsigned int* n_events, cudaStream_t st);

}  // namespace host
}  // namespace b200vq

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

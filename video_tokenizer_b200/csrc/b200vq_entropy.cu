// b200vq -- C ABI (include/b200vq.h) over the sm_100a kernels.  No torch types here.
// This unit: the streaming entropy regulariser.
#include "vq_entropy.cuh"
#include "b200vq_host.h"

using namespace b200vq;
using namespace b200vq::host;

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

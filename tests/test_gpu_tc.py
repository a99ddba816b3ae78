"""GPU: the tcgen05 tensor-core scoring path against the exact-fp32 CUDA-core path and the
oracle.  Deterministic modes must be BIT-identical to the exact path (the tensor-core pass
only decides rows whose approximate top-2 gap exceeds twice its error bound; the rest are
re-scored in fp32), stochastic mode must pick the same code as the exact Gumbel-max except
where two perturbed scores are within the split-bf16 error."""
import pytest
import torch

import video_tokenizer_b200 as V
from video_tokenizer_b200 import _lib
from video_tokenizer_b200 import functional as VF
from oracle import vq_oracle as O
from tests.helpers import assert_indices_match

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _scores(z, w, mode, tc, seed=7, offset=12, t_inv=1 / 0.03, row_offset=0):
    lib = _lib.load()
    N, D = z.shape
    K = w.shape[0]
    emb, e2, _ = VF.prepare_codebook(w, True)
    keys = torch.zeros(N, dtype=torch.int64, device=DEV)
    nres = torch.zeros(1, dtype=torch.int32, device=DEV)
    dt = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}[z.dtype]
    if tc:
        image = VF.prepare_codebook_tc(emb)
        second = torch.zeros(2 * N + 4, dtype=torch.float32, device=DEV)
        _lib.check(lib.b200vq_scores_tc(z.data_ptr(), dt, D, N, D, image.data_ptr(), emb.data_ptr(), e2.data_ptr(), K, 0,
                                        mode, 1, t_inv, seed, offset, row_offset, keys.data_ptr(), second.data_ptr(),
                                        nres.data_ptr(), _stream()), "scores_tc")
    else:
        _lib.check(lib.b200vq_scores_partial(z.data_ptr(), dt, D, N, D, emb.data_ptr(), e2.data_ptr(), K, 0, mode, 1,
                                             t_inv, seed, offset, row_offset, keys.data_ptr(), _stream()), "scores")
    torch.cuda.synchronize()
    return (~keys) & 0xFFFFFFFF, int(nres.item())


@pytest.mark.parametrize("D,K,N", [(24, 8192, 4096), (16, 8192, 4096), (16, 500, 1000), (8, 64, 130), (32, 777, 257),
                                   (24, 256, 1), (16, 1024, 128), (24, 1000, 129)])
@pytest.mark.parametrize("mode", [_lib.MODE_ARGMIN_DIST, _lib.MODE_ARGMAX_COS])
def test_tc_deterministic_equals_exact(D, K, N, mode):
    torch.manual_seed(D * 1000 + K)
    z = torch.randn(N, D, device=DEV)
    w = torch.randn(K, D, device=DEV)
    exact, _ = _scores(z, w, mode, tc=False)
    tc, nres = _scores(z, w, mode, tc=True)
    assert torch.equal(tc, exact), f"{int((tc != exact).sum())} of {N} rows differ"
    assert nres <= max(8, N // 20), f"{nres} of {N} rows needed exact re-scoring"


def test_tc_edge_rows_ties_zero_and_clusters():
    """duplicate codes (exact ties -> lowest index), all-zero latent rows, a zero code row,
    latents sitting exactly on codes, tightly clustered codes."""
    torch.manual_seed(5)
    D, K, N = 16, 1024, 640
    w = torch.randn(K, D, device=DEV)
    w[100] = w[7]
    w[900] = w[7]
    w[33] = 0
    w[500:564] = w[400] + 1e-4 * torch.randn(64, D, device=DEV)
    z = torch.randn(N, D, device=DEV)
    z[0] = 0
    z[1] = w[7] * 3
    z[2] = w[400]
    z[3:67] = w[500:564]
    for mode in (_lib.MODE_ARGMIN_DIST, _lib.MODE_ARGMAX_COS):
        exact, _ = _scores(z, w, mode, tc=False)
        tc, nres = _scores(z, w, mode, tc=True)
        assert torch.equal(tc, exact)
        assert int(tc[1]) == 7
        assert nres >= 60


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_tc_half_inputs(dtype):
    torch.manual_seed(6)
    z = torch.randn(2000, 24, device=DEV).to(dtype)
    w = torch.randn(2048, 24, device=DEV)
    exact, _ = _scores(z, w, _lib.MODE_ARGMIN_DIST, tc=False)
    tc, _ = _scores(z, w, _lib.MODE_ARGMIN_DIST, tc=True)
    assert torch.equal(tc, exact)


def test_tc_full_size_vs_exact_and_oracle():
    """BASELINE size (64 x 1024 tokens, K = 8192, D = 24): identical to the exact path on every
    row; oracle parity on a bounded sample."""
    torch.manual_seed(8)
    N, D, K = 65536, 24, 8192
    z = torch.randn(N, D, device=DEV)
    w = torch.empty(K, D, device=DEV)
    torch.nn.init.kaiming_uniform_(w)
    exact, _ = _scores(z, w, _lib.MODE_ARGMIN_DIST, tc=False)
    tc, nres = _scores(z, w, _lib.MODE_ARGMIN_DIST, tc=True)
    assert torch.equal(tc, exact)
    assert nres < N // 50, nres
    rows = torch.randperm(N)[:1024]
    cfg = O.VQConfig(dim=D, codebook_size=K, l2_normalized=True)
    ref = O.vq_forward(z[rows].cpu().reshape(1, -1, D), w.cpu(), cfg, return_scores=True)
    assert_indices_match(tc[rows], ref["bottleneck_rep"], ref["scores"], False)


@pytest.mark.parametrize("D,K,N", [(24, 8192, 4096), (16, 1000, 777), (8, 256, 128), (32, 5000, 300), (24, 65536, 2048),
                                   (16, 300, 1)])
def test_tc_sampling_matches_exact_gumbel_max(D, K, N):
    """Same Philox stream, same per-(row, code) noise: the lazy tensor-core sampler only FILTERS with its bf16
    scores; every candidate is expanded in exact fp32 with the arithmetic of the exact kernel, so the draw is
    the exact kernel's draw on every row."""
    torch.manual_seed(9)
    z = torch.randn(N, D, device=DEV)
    w = torch.randn(K, D, device=DEV)
    exact, _ = _scores(z, w, _lib.MODE_SAMPLE, tc=False, seed=3, offset=40, row_offset=5)
    tc, _ = _scores(z, w, _lib.MODE_SAMPLE, tc=True, seed=3, offset=40, row_offset=5)
    assert torch.equal(tc, exact), f"{int((tc != exact).sum())} of {N} sampled codes differ"
    other, _ = _scores(z, w, _lib.MODE_SAMPLE, tc=True, seed=3, offset=44, row_offset=5)
    if N >= 64:
        assert int((other != tc).sum()) > N // 4   # a new offset is a new draw


def test_module_uses_tc_by_default_and_matches_exact():
    torch.manual_seed(10)
    m = V.SimpleVectorQuantizer(dim=24, codebook_size=8192, l2_normalized=True).to(DEV)
    z = torch.randn(8, 1024, 24, device=DEV, requires_grad=True)
    assert VF.SCORING == "tc"
    a = m(z)
    prev = VF.set_scoring("exact")
    try:
        b = m(z)
    finally:
        VF.set_scoring(prev)
    assert torch.equal(a["bottleneck_rep"], b["bottleneck_rep"])
    assert torch.equal(a["regularized_z"], b["regularized_z"])
    assert float(a["loss_q"]) == float(b["loss_q"])


@pytest.mark.parametrize("t_inv", [1.0, 10.0, 300.0, 3000.0])
def test_tc_sampling_equals_exact_at_any_temperature(t_inv):
    """the filter's slack scales with 1/T (score error x t_inv): draws stay the exact kernel's at any temperature"""
    torch.manual_seed(11)
    z = torch.randn(3000, 24, device=DEV)
    w = torch.randn(4096, 24, device=DEV)
    exact, _ = _scores(z, w, _lib.MODE_SAMPLE, tc=False, seed=5, offset=8, t_inv=t_inv)
    tc, _ = _scores(z, w, _lib.MODE_SAMPLE, tc=True, seed=5, offset=8, t_inv=t_inv)
    assert torch.equal(tc, exact), f"{int((tc != exact).sum())} rows differ at t_inv={t_inv}"


def test_tc_sampling_full_size_equals_exact_and_is_repeatable():
    """BASELINE size, stochastic: equal to the exact kernel on all 65 536 rows, and identical keys over repeated
    launches (the only race detector available: the kernel has mbarrier rings, a shared-memory queue and atomics)."""
    torch.manual_seed(12)
    N, D, K = 65536, 24, 8192
    z = torch.randn(N, D, device=DEV)
    w = torch.empty(K, D, device=DEV)
    torch.nn.init.kaiming_uniform_(w)
    exact, _ = _scores(z, w, _lib.MODE_SAMPLE, tc=False, seed=21, offset=4)
    first, _ = _scores(z, w, _lib.MODE_SAMPLE, tc=True, seed=21, offset=4)
    assert torch.equal(first, exact), f"{int((first != exact).sum())} of {N} rows differ"
    for _ in range(9):
        again, _ = _scores(z, w, _lib.MODE_SAMPLE, tc=True, seed=21, offset=4)
        assert torch.equal(again, first)


@pytest.mark.parametrize("mode", [_lib.MODE_ARGMIN_DIST, _lib.MODE_ARGMAX_COS])
def test_tc_deterministic_full_size_repeatable(mode):
    """same keys over 10 launches at the BASELINE size in both deterministic flavours"""
    torch.manual_seed(13)
    N, D, K = 65536, 24, 8192
    z = torch.randn(N, D, device=DEV)
    w = torch.empty(K, D, device=DEV)
    torch.nn.init.kaiming_uniform_(w)
    first, _ = _scores(z, w, mode, tc=True)
    for _ in range(9):
        again, _ = _scores(z, w, mode, tc=True)
        assert torch.equal(again, first)

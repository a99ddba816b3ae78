"""b200vq_weight_grad_noalloc: the allocation-free variant of b200vq_weight_grad_ws (scratch = a static device array
inside the library, include/b200vq.h).  Collected last on purpose (pytest -x): a failure here hides no other test.

No counterpart in the reference; the comparison is against the library's own workspace entry point.
"""
import pytest
import torch

import video_tokenizer_b200 as V
from video_tokenizer_b200 import _lib

ENOALLOC_SIZE = -6
LIMIT = 24 << 20


def rel_err(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def test_sizes_beyond_the_built_in_scratch_are_refused_before_any_cuda_work():
    """host-side argument checks only: runs without a GPU"""
    lib = V.load_library()
    assert int(lib.b200vq_weight_grad_workspace_bytes(1024, 32)) > LIMIT
    assert lib.b200vq_weight_grad_noalloc(None, 1024, None, 4096, 1024, 32, None, 0, None, None) == ENOALLOC_SIZE
    assert b"built-in scratch" in lib.b200vq_error_string(ENOALLOC_SIZE)
    assert lib.b200vq_weight_grad_noalloc(None, 768, None, -1, 768, 24, None, 0, None, None) == -1      # N < 0
    assert lib.b200vq_weight_grad_noalloc(None, 770, None, 16, 770, 24, None, 0, None, None) == -2      # C % 16
    with pytest.raises(_lib.B200VQError, match="built-in scratch"):
        _lib.check(ENOALLOC_SIZE, "b200vq_weight_grad_noalloc")


def _both(lib, x, y, C, D, dc, ws):
    """(dw, colsum) from the workspace entry point and from the allocation-free one, same inputs"""
    N = x.shape[0]
    st = torch.cuda.current_stream().cuda_stream
    shape = (D, C) if dc else (C, D)
    dw_ws, cs_ws = torch.full(shape, 7.0, device="cuda"), torch.full((C,), 7.0, device="cuda")
    dw_na, cs_na = torch.full(shape, -3.0, device="cuda"), torch.full((C,), -3.0, device="cuda")
    _lib.check(lib.b200vq_weight_grad_ws(x.data_ptr(), C, y.data_ptr(), N, C, D, dw_ws.data_ptr(), dc, cs_ws.data_ptr(),
                                         ws.data_ptr(), ws.numel(), st), "b200vq_weight_grad_ws")
    before = torch.cuda.memory_allocated()
    _lib.check(lib.b200vq_weight_grad_noalloc(x.data_ptr(), C, y.data_ptr(), N, C, D, dw_na.data_ptr(), dc, cs_na.data_ptr(), st),
               "b200vq_weight_grad_noalloc")
    assert torch.cuda.memory_allocated() == before
    torch.cuda.synchronize()
    return (dw_ws, cs_ws), (dw_na, cs_na)


@pytest.mark.gpu
@pytest.mark.parametrize("C,D,dc", [(768, 24, 0), (528, 32, 1)])
def test_noalloc_weight_grad_matches_the_workspace_entry_point_on_two_calls_in_a_row(C, D, dc):
    """two calls in a row with different inputs and sizes: the second one starts from the first one's slabs.
    (768, 24) takes the token-outer kernel (one slab per SM), (528, 32) the chunk-outer one (two per SM);
    the small-N call falls back to atomics into zeroed outputs."""
    lib = V.load_library()
    assert int(lib.b200vq_weight_grad_workspace_bytes(C, D)) <= LIMIT
    ws = torch.empty(int(lib.b200vq_weight_grad_workspace_bytes(C, D)), dtype=torch.uint8, device="cuda")
    torch.manual_seed(C + D)
    for call, N in enumerate((40000, 65536, 3000)):
        x = torch.randn(N, C, device="cuda") * (call + 1)
        y = torch.randn(N, D, device="cuda")
        ws.fill_(0xFF)
        (dw_ws, cs_ws), (dw_na, cs_na) = _both(lib, x, y, C, D, dc, ws)
        ref = x.double().t() @ y.double()
        ref = ref.t() if dc else ref
        assert rel_err(dw_na, ref) < 1e-3 and rel_err(cs_na, x.double().sum(0)) < 1e-5
        assert rel_err(dw_na, dw_ws) < 1e-3 and rel_err(cs_na, cs_ws) < 1e-5
        if N >= 296 * 64:   # slab path: same kernels, same slab order -> the same bits
            assert torch.equal(dw_na, dw_ws) and torch.equal(cs_na, cs_ws)
    # N == 0: outputs are zeroed, nothing is launched
    x0, y0 = torch.empty(0, C, device="cuda"), torch.empty(0, D, device="cuda")
    (dw_ws, cs_ws), (dw_na, cs_na) = _both(lib, x0, y0, C, D, dc, ws)
    assert dw_na.abs().max() == 0 and cs_na.abs().max() == 0 and torch.equal(dw_na, dw_ws)


@pytest.mark.gpu
def test_noalloc_weight_grad_under_stream_capture():
    """the call is legal inside a CUDA graph capture (an allocation would invalidate the capture) and the replays
    reproduce the eager result"""
    lib = V.load_library()
    C, D, N = 768, 24, 32768
    torch.manual_seed(5)
    x, y = torch.randn(N, C, device="cuda"), torch.randn(N, D, device="cuda")
    dw, cs = torch.zeros(C, D, device="cuda"), torch.zeros(C, device="cuda")
    dw_g, cs_g = torch.zeros(C, D, device="cuda"), torch.zeros(C, device="cuda")

    def call(dw_, cs_):
        _lib.check(lib.b200vq_weight_grad_noalloc(x.data_ptr(), C, y.data_ptr(), N, C, D, dw_.data_ptr(), 0, cs_.data_ptr(),
                                                  torch.cuda.current_stream().cuda_stream), "b200vq_weight_grad_noalloc")

    call(dw, cs)                       # first call on this device: resolves the scratch address outside the capture
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        call(dw_g, cs_g)
    for _ in range(2):
        dw_g.fill_(1.0)
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(dw_g, dw) and torch.equal(cs_g, cs)

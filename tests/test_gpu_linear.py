"""GPU: the streaming in/out linear kernels (include/b200vq.h: b200vq_linear_*) against fp64
references.  in_linear must be fp32-class (it decides code indices); the others must be inside
the 1e-3 relative tolerance of BASELINE's north_star."""
import pytest
import torch

import video_tokenizer_b200 as V
from video_tokenizer_b200 import functional as VF
from tests.helpers import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("N,C,D", [(4096, 768, 24), (1000, 768, 16), (129, 512, 8), (1, 64, 32), (300, 1024, 24),
                                   (40000, 768, 24), (19001, 272, 16)])
def test_linear_in_out_forward_backward(N, C, D):
    torch.manual_seed(N + C + D)
    x = torch.randn(N, C, device=DEV, requires_grad=True)
    w_in = (torch.randn(D, C, device=DEV) / C ** 0.5).requires_grad_(True)
    b_in = torch.randn(D, device=DEV, requires_grad=True)
    w_out = (torch.randn(C, D, device=DEV) / D ** 0.5).requires_grad_(True)
    b_out = torch.randn(C, device=DEV, requires_grad=True)
    assert VF.linears_supported(x, w_in, C, D)
    z = VF.linear_in(x, w_in, b_in)
    y = VF.linear_out(z, w_out, b_out)
    G = torch.randn(N, C, device=DEV)
    (y * G).sum().backward()
    # fp64 reference
    xd, wid, bid, wod, bod = [t.detach().double().requires_grad_(True) for t in (x, w_in, b_in, w_out, b_out)]
    zd = torch.nn.functional.linear(xd, wid, bid)
    yd = torch.nn.functional.linear(zd, wod, bod)
    (yd * G.double()).sum().backward()
    assert rel_err(z, zd) < 2e-6, rel_err(z, zd)
    assert (z.detach().double() - zd).abs().max() / zd.abs().max() < 5e-6
    assert rel_err(y, yd) < 1e-3
    for name, a, b in (("grad_x", x.grad, xd.grad), ("grad_w_in", w_in.grad, wid.grad), ("grad_b_in", b_in.grad, bid.grad),
                       ("grad_w_out", w_out.grad, wod.grad), ("grad_b_out", b_out.grad, bod.grad)):
        assert rel_err(a, b) < 1e-3, (name, rel_err(a, b))


@pytest.mark.parametrize("N,C,D,dc", [(65536, 768, 24, 0), (30011, 528, 32, 1), (5000, 768, 24, 0), (0, 768, 24, 0)])
def test_weight_grad_ws_matches_fp64_and_is_deterministic(N, C, D, dc):
    """b200vq_weight_grad_ws: per-CTA slabs + ordered reduction (no atomics); small N falls back to atomics"""
    from video_tokenizer_b200 import _lib
    lib = V.load_library()
    torch.manual_seed(N + C)
    x = torch.randn(N, C, device=DEV)
    y = torch.randn(N, D, device=DEV)
    ws = torch.empty(int(lib.b200vq_weight_grad_workspace_bytes(C, D)), dtype=torch.uint8, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    outs = []
    for _ in range(2):
        ws.fill_(0xFF)                                # NaN patterns: every slab that is read must have been written
        dw = torch.full((D, C) if dc else (C, D), 7.0, device=DEV)
        cs = torch.full((C,), 7.0, device=DEV)
        _lib.check(lib.b200vq_weight_grad_ws(x.data_ptr(), C, y.data_ptr(), N, C, D, dw.data_ptr(), dc, cs.data_ptr(),
                                             ws.data_ptr(), ws.numel(), st), "wg_ws")
        outs.append((dw, cs))
    ref = x.double().t() @ y.double()
    ref = ref.t() if dc else ref
    if N == 0:
        assert outs[0][0].abs().max() == 0 and outs[0][1].abs().max() == 0
        return
    assert rel_err(outs[0][0], ref) < 1e-3
    assert rel_err(outs[0][1], x.double().sum(0)) < 1e-5
    if N >= 296 * 64:
        assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert lib.b200vq_weight_grad_ws(x.data_ptr(), C, y.data_ptr(), N, C, D, dw.data_ptr(), dc, cs.data_ptr(),
                                     ws.data_ptr(), 16, st) != 0       # workspace too small


@pytest.mark.parametrize("N,C,D,add", [(65536, 768, 24, True), (4099, 512, 16, False), (31, 256, 8, True), (1, 1024, 24, False),
                                        (10000, 512, 32, True), (33, 768, 24, False)])
def test_grad_out_single_pass_matches_fp64(N, C, D, add):
    """b200vq_grad_out: grad of out_linear's input, weight and bias from one pass over grad_output"""
    from video_tokenizer_b200 import _lib
    lib = V.load_library()
    assert lib.b200vq_grad_out_supported(C, D)
    torch.manual_seed(N + C + D)
    big = torch.randn(N, C + 8, device=DEV)
    g = big[:, :C]                                        # row stride C + 8
    a = torch.randn(N, D, device=DEV)
    w = torch.randn(C, D, device=DEV) / D ** 0.5
    extra = torch.randn(N, D, device=DEV) if add else None
    ga = torch.full((N, D), 9.0, device=DEV); dw = torch.full((C, D), 9.0, device=DEV); db = torch.full((C,), 9.0, device=DEV)
    ws = torch.empty(int(lib.b200vq_weight_grad_workspace_bytes(C, D)), dtype=torch.uint8, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    outs = []
    for _ in range(2):
        ws.fill_(0xFF)
        _lib.check(lib.b200vq_grad_out(g.data_ptr(), C + 8, a.data_ptr(), w.data_ptr(), N, C, D, extra.data_ptr() if add else None,
                                       ga.data_ptr(), dw.data_ptr(), db.data_ptr(), ws.data_ptr(), ws.numel(), st), "grad_out")
        outs.append((ga.clone(), dw.clone(), db.clone()))
    gd = g.double()
    ref_ga = gd @ w.double() + (extra.double() if add else 0)
    assert rel_err(outs[0][0], ref_ga) < 1e-3, rel_err(outs[0][0], ref_ga)
    assert rel_err(outs[0][1], gd.t() @ a.double()) < 1e-3
    assert rel_err(outs[0][2], gd.sum(0)) < 1e-5
    for x, y in zip(outs[0], outs[1]):
        assert torch.equal(x, y)                          # no atomics: run-to-run identical
    assert not lib.b200vq_grad_out_supported(272, 16) and not lib.b200vq_grad_out_supported(1280, 24)


def test_linear_in_strided_rows_and_3d():
    """x is a batch-strided slice of a larger activation (models/transformer.py:69)."""
    torch.manual_seed(0)
    B, n, C, D = 3, 100, 256, 16
    big = torch.randn(B, 2 * n, C, device=DEV)
    x = big[:, -n:, :]
    assert not x.is_contiguous()
    w = torch.randn(D, C, device=DEV) / 16
    b = torch.randn(D, device=DEV)
    z = VF.linear_in(x, w, b)
    ref = torch.nn.functional.linear(x.double(), w.double(), b.double())
    assert z.shape == (B, n, D)
    assert rel_err(z, ref) < 2e-6


def test_bottleneck_fused_linears_match_torch_linears():
    torch.manual_seed(1)
    spec = {"name": "bottleneck", "args": {"bottleneck_dim": 24, "norm": "none", "regularizer": {"name": "vq", "args": {
        "codebook_size": 1024, "l2_normalized": True, "stochastic": False}}}}
    m = V.make(spec, args={"token_nums": 64, "input_dim": 768, "output_dim": 768}).to(DEV)
    x = torch.randn(4, 64, 768, device=DEV)
    outs = {}
    for kind in ("fused", "torch"):
        prev = VF.set_linears(kind)
        try:
            xr = x.clone().requires_grad_(True)
            for p in m.parameters():
                p.grad = None
            o = m(xr)
            (o["output"].square().mean() + 0.1 * o["loss_q"]).backward()
            outs[kind] = (o, xr.grad.clone(), {k: p.grad.clone() for k, p in m.named_parameters()})
        finally:
            VF.set_linears(prev)
    of, gxf, gpf = outs["fused"]
    ot, gxt, gpt = outs["torch"]
    same = (of["bottleneck_rep"] == ot["bottleneck_rep"]).float().mean().item()
    assert same >= 0.995, same
    if same == 1.0:
        assert rel_err(of["output"], ot["output"]) < 1e-3
        assert rel_err(gxf, gxt) < 2e-3
        for k in gpf:
            assert rel_err(gpf[k], gpt[k]) < 2e-3, k

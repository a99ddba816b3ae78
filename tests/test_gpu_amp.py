"""The drop-in `Bottleneck` under autocast (BASELINE configs[3]: the bf16 tokenizer step; SURVEY.md 8b "dtype flow
under AMP"; reference models/bottleneck.py:140-188 with the quantizer's autocast-off fp32 island :250-264).

The linears follow the caller's autocast dtype exactly like the reference's `nn.Linear`s (projected_z / output come
back bf16 or fp16), the quantizer runs in fp32 on `z.float()` and hands fp32 tensors back, indices are int64.  Parity
is judged on identical latents: the oracle quantizer is fed OUR projected_z, and the oracle's whole-module backward is
run under CPU autocast with the indices forced to ours."""
import pytest
import torch
import torch.nn.functional as F

from oracle import vq_oracle as O
from tests.helpers import assert_close, assert_indices_match

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _module(D, K, C, n, stochastic):
    import video_tokenizer_b200 as V
    spec = {"name": "bottleneck", "args": {"bottleneck_dim": D, "norm": "none", "regularizer": {"name": "vq", "args": {
        "codebook_size": K, "commitment_loss_weight": 0.25, "codebook_loss_weight": 1.0, "entropy_loss_weight": 0.0,
        "entropy_loss_temperature": 0.01, "l2_normalized": True, "stochastic": stochastic,
        "stochastic_temperature": 0.03}}}}
    return V.make(spec, args={"token_nums": n, "input_dim": C, "output_dim": C}).to(DEV)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("stochastic", [False, True])
def test_dtype_flow_and_quantizer_parity_under_autocast(dtype, stochastic):
    torch.manual_seed(7)
    B, n, C, D, K = 3, 128, 768, 16, 8192          # the larp_tokenizer_large.yaml bottleneck block (:39-54)
    m = _module(D, K, C, n, stochastic)
    if stochastic:                                   # deterministic flavour of the stochastic config (:277-278)
        m.eval()
        m.regularizer.set_eval_deterministic(True)
    big = torch.randn(B, 2 * n, C, device=DEV)
    x = big[:, -n:, :].detach().requires_grad_(True)   # batch-strided slice, fp32 residual stream
    with torch.autocast("cuda", dtype=dtype):
        out = m(x)
    assert out["output"].dtype == dtype and out["projected_z"].dtype == dtype
    for k in ("regularized_z", "unregularized_z", "emb", "loss_q", "loss_commit", "loss_codebook"):
        assert out[k].dtype == torch.float32, k
    assert out["bottleneck_rep"].dtype == torch.int64 and out["bottleneck_rep"].shape == (B, n)
    # the linears are the caller's autocast linears
    with torch.autocast("cuda", dtype=dtype):
        z_ref = F.linear(x, m.in_linear.weight, m.in_linear.bias)
    assert torch.equal(out["projected_z"], z_ref)
    # quantizer on identical (bf16 / fp16) latents against the oracle
    W = m.regularizer.embedding.weight.detach().cpu()
    cfg = O.VQConfig(dim=D, codebook_size=K, l2_normalized=True, stochastic=stochastic, stochastic_temperature=0.03)
    ref = O.vq_forward(out["projected_z"].detach().cpu(), W, cfg, training=not stochastic,
                       eval_deterministic=stochastic, return_scores=True)
    assert_indices_match(out["bottleneck_rep"], ref["bottleneck_rep"], ref["scores"], stochastic)
    same = bool((out["bottleneck_rep"].cpu() == ref["bottleneck_rep"]).all())
    for k in ("regularized_z", "unregularized_z", "loss_q", "loss_commit", "loss_codebook"):
        assert_close(out[k].detach().cpu(), ref[k].detach(), k, 1e-3 if same else 5e-3)
    with torch.autocast("cuda", dtype=dtype):
        o_ref = F.linear(out["regularized_z"].detach(), m.out_linear.weight, m.out_linear.bias)
    assert torch.equal(out["output"].detach(), o_ref)


def test_backward_under_bf16_autocast_against_oracle():
    """whole module fwd+bwd under bf16 autocast vs the oracle under CPU bf16 autocast, indices forced to ours.
    Tolerance: the two bf16 GEMM implementations (cuBLAS vs oneDNN) round differently -> 1e-2 on tensors that pass
    through a bf16 product, 1e-3 on the fp32 quantizer tensors."""
    torch.manual_seed(11)
    B, n, C, D, K = 2, 64, 256, 16, 1024
    m = _module(D, K, C, n, False)
    x = torch.randn(B, n, C, device=DEV, requires_grad=True)
    G = torch.randn(B, n, C, device=DEV)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = m(x)
        obj = (out["output"].float() * G).sum() + 0.1 * out["loss_q"]
    obj.backward()
    p = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    xr = x.detach().cpu().clone().requires_grad_(True)
    cfg = O.VQConfig(dim=D, codebook_size=K, l2_normalized=True)
    with torch.autocast("cpu", dtype=torch.bfloat16):
        z = F.linear(xr, p["in_linear.weight"], p["in_linear.bias"])
    # the quantizer is an autocast-off fp32 island (bottleneck.py:250, :260-264)
    r = O.vq_forward(z, p["regularizer.embedding.weight"], cfg, forced_indices=out["bottleneck_rep"].cpu().reshape(-1))
    with torch.autocast("cpu", dtype=torch.bfloat16):
        o = F.linear(r["regularized_z"], p["out_linear.weight"], p["out_linear.bias"])
    robj = (o.float() * G.cpu()).sum() + 0.1 * r["loss_q"]
    robj.backward()
    assert o.dtype == torch.bfloat16 and z.dtype == torch.bfloat16
    assert_close(out["output"].float().cpu(), o.float(), "output", 1e-2)
    assert_close(out["loss_q"].detach().cpu(), r["loss_q"].detach(), "loss_q", 1e-2)
    assert_close(x.grad.cpu(), xr.grad, "grad_x", 2e-2)
    assert_close(m.in_linear.weight.grad.cpu(), p["in_linear.weight"].grad, "grad_w_in", 2e-2)
    assert_close(m.out_linear.weight.grad.cpu(), p["out_linear.weight"].grad, "grad_w_out", 2e-2)
    assert_close(m.out_linear.bias.grad.cpu(), p["out_linear.bias"].grad, "grad_b_out", 2e-2)
    assert_close(m.regularizer.embedding.weight.grad.cpu(), p["regularizer.embedding.weight"].grad, "grad_codebook", 2e-2)


def test_backward_is_linear_in_the_loss_scale():
    """fp16 + GradScaler (scripts/train_larp_tokenizer_reproduce.sh:30, trainers/larp_tokenizer_trainer.py:218-227):
    incoming gradients are scaled, so the fused backward must be linear in them"""
    torch.manual_seed(13)
    m = _module(16, 512, 256, 32, False)
    x = torch.randn(2, 32, 256, device=DEV)
    G = torch.randn(2, 32, 256, device=DEV)
    grads = []
    for scale in (1.0, 1024.0):
        m.zero_grad(set_to_none=True)
        xs = x.clone().requires_grad_(True)
        with torch.autocast("cuda", dtype=torch.float16):
            out = m(xs)
            obj = ((out["output"].float() * G).sum() + 0.1 * out["loss_q"]) * scale
        obj.backward()
        grads.append([xs.grad.clone()] + [p.grad.clone() for p in m.parameters()])
    for a, b in zip(*grads):
        assert_close(b / 1024.0, a, "scaled grads", 2e-3)

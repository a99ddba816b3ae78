"""The drop-in `Bottleneck` under autocast (BASELINE configs[3]: the bf16 tokenizer step; SURVEY.md 8b "dtype flow
under AMP"; reference models/bottleneck.py:140-188 with the quantizer's autocast-off fp32 island :250-264).

The linears follow the caller's autocast dtype like the reference's `nn.Linear`s (projected_z / output come back bf16
or fp16) but run on the library's own streaming kernels -- no cuBLAS launch in the step: fp32 x and weights as TF32
operands, fp32 accumulation, ONE rounding to the 2-byte type (the reference's autocast GEMM rounds x and W to bf16 first,
so the two results differ by at most a couple of units in the last place of the 2-byte output).  The quantizer runs in
fp32 on `z.float()` and hands fp32 tensors back, indices are int64.  Parity is judged on identical latents: the oracle
quantizer is fed OUR projected_z, and the oracle's whole-module backward is run under CPU autocast with the indices
forced to ours."""
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
    # in_linear: against the fp32 product rounded once (tight), and against the reference's autocast nn.Linear (cuBLAS,
    # inputs rounded to the 2-byte type first: a couple of ulps of the 2-byte output away)
    ulp = 2.0 ** -8 if dtype is torch.bfloat16 else 2.0 ** -11
    z32 = F.linear(x.detach(), m.in_linear.weight, m.in_linear.bias)
    assert_close(out["projected_z"].float(), z32, "projected_z vs fp32", ulp)
    with torch.autocast("cuda", dtype=dtype):
        z_ref = F.linear(x, m.in_linear.weight, m.in_linear.bias)
    assert_close(out["projected_z"].float(), z_ref.float(), "projected_z vs autocast nn.Linear", 2.5 * 2.0 ** -8)
    # quantizer on identical (bf16 / fp16) latents against the oracle
    W = m.regularizer.embedding.weight.detach().cpu()
    cfg = O.VQConfig(dim=D, codebook_size=K, l2_normalized=True, stochastic=stochastic, stochastic_temperature=0.03)
    ref = O.vq_forward(out["projected_z"].detach().cpu(), W, cfg, training=not stochastic,
                       eval_deterministic=stochastic, return_scores=True)
    assert_indices_match(out["bottleneck_rep"], ref["bottleneck_rep"], ref["scores"], stochastic)
    same = bool((out["bottleneck_rep"].cpu() == ref["bottleneck_rep"]).all())
    for k in ("regularized_z", "unregularized_z", "loss_q", "loss_commit", "loss_codebook"):
        assert_close(out[k].detach().cpu(), ref[k].detach(), k, 1e-3 if same else 5e-3)
    o32 = F.linear(out["regularized_z"].detach(), m.out_linear.weight, m.out_linear.bias)
    assert_close(out["output"].detach().float(), o32, "output vs fp32", ulp)
    with torch.autocast("cuda", dtype=dtype):
        o_ref = F.linear(out["regularized_z"].detach(), m.out_linear.weight, m.out_linear.bias)
    assert_close(out["output"].detach().float(), o_ref.float(), "output vs autocast nn.Linear", 2.5 * 2.0 ** -8)


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


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_one_call_path_under_autocast_launches_no_library_gemm(dtype):
    """BASELINE configs[3]: under autocast the whole step stays on the extension's kernels (functional.bottleneck_supported
    is true, two host calls per step) -- the profiler sees no cuBLAS / CUTLASS GEMM kernel at all"""
    from video_tokenizer_b200 import functional as VF
    torch.manual_seed(17)
    B, n, C, D, K = 4, 1024, 768, 16, 8192
    m = _module(D, K, C, n, True)
    m.train()
    x = torch.randn(B, n, C, device=DEV, requires_grad=True)
    G = torch.randn(B, n, C, device=DEV)
    with torch.autocast("cuda", dtype=dtype):
        assert VF.bottleneck_supported(x, m.in_linear.weight, m.regularizer.embedding.weight, m.out_linear.weight)

    def step():
        with torch.autocast("cuda", dtype=dtype):
            out = m(x)
            obj = (out["output"].float() * G).sum() + 0.1 * out["loss_q"]
        obj.backward()
        return out

    step()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        out = step()
        torch.cuda.synchronize()
    names = [e.key for e in prof.key_averages()]
    gemms = [k for k in names if any(t in k.lower() for t in ("gemm", "cutlass", "cublas", "sgemm", "xmma", "nvjet"))]
    assert not gemms, gemms
    assert any("linear_in_kernel" in k for k in names) and any("grad_out_kernel" in k for k in names), names
    assert out["output"].dtype == dtype and x.grad.dtype == torch.float32
    assert all(p.grad is not None and p.grad.dtype == torch.float32 and torch.isfinite(p.grad).all() for p in m.parameters())


def test_training_forward_folds_the_decoder_positional_table():
    """`z + decoder_latent_pe` (models/larp_tokenizer.py:463-464) folded into the out_linear store of
    the TRAINING forward: Bottleneck.forward(x, add=pe) == Bottleneck.forward(x)['output'] + pe, same gradients (the
    table is a buffer in every shipped config: learned_decoder_latent_pe = False)"""
    torch.manual_seed(19)
    B, n, C, D, K = 3, 256, 768, 24, 4096
    m = _module(D, K, C, n, False)
    x = torch.randn(B, n, C, device=DEV)
    pe = torch.randn(1, n, C, device=DEV)
    G = torch.randn(B, n, C, device=DEV)
    res = []
    for fold in (False, True):
        m.zero_grad(set_to_none=True)
        xs = x.clone().requires_grad_(True)
        out = m(xs, add=pe[0]) if fold else m(xs)
        y = out["output"] if fold else out["output"] + pe
        ((y * G).sum() + 0.1 * out["loss_q"]).backward()
        res.append((y.detach(), xs.grad.clone(), [p.grad.clone() for p in m.parameters()], out["bottleneck_rep"]))
    assert torch.equal(res[0][3], res[1][3])
    assert_close(res[1][0], res[0][0], "output + pe", 1e-6)
    assert_close(res[1][1], res[0][1], "grad_x", 1e-6)
    for a, b in zip(res[1][2], res[0][2]):
        assert_close(a, b, "parameter grads", 1e-6)
    # against the oracle as well
    p = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    cfg = O.VQConfig(dim=D, codebook_size=K, l2_normalized=True)
    ref = O.bottleneck_forward(x.cpu(), p["in_linear.weight"], p["in_linear.bias"], p["out_linear.weight"], p["out_linear.bias"],
                               p["regularizer.embedding.weight"], cfg, training=True)
    if torch.equal(ref["bottleneck_rep"].reshape(-1), res[1][3].cpu().reshape(-1)):
        assert_close(res[1][0].cpu(), ref["output"] + pe.cpu(), "output + pe vs oracle")

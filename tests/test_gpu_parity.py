"""GPU parity: the CUDA path (through the C ABI and the drop-in modules) against the
oracle and the reference-generated golden vectors.  Tolerances are BASELINE.json's:
indices bit-exact outside a 1e-6 reference score gap; tensors/losses/grads 1e-3 relative."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import video_tokenizer_b200 as V
from oracle import vq_oracle as O
from tests.helpers import (assert_close, assert_indices_match, cfg_from_golden, golden_z, load_golden,
                           vq_golden_names)

pytestmark = pytest.mark.gpu
DEV = "cuda"


def make_vq(cfg: O.VQConfig, weight, eval_det=False):
    m = V.SimpleVectorQuantizer(dim=cfg.dim, codebook_size=cfg.codebook_size,
                                commitment_loss_weight=cfg.commitment_loss_weight,
                                codebook_loss_weight=cfg.codebook_loss_weight, l2_normalized=cfg.l2_normalized,
                                same_index_shape=cfg.same_index_shape, stochastic=cfg.stochastic,
                                stochastic_temperature=cfg.stochastic_temperature).to(DEV)
    with torch.no_grad():
        m.embedding.weight.copy_(torch.as_tensor(weight))
    if eval_det:
        m.eval()
        m.set_eval_deterministic(True)
    return m


def run_and_compare(z_cpu, w_cpu, cfg, eval_det, G=None, g_lossq=0.7, tol=1e-3):
    """CUDA module vs oracle on identical latents; returns number of (allowed) index mismatches."""
    zr = z_cpu.clone().requires_grad_(True)
    wr = w_cpu.clone().requires_grad_(True)
    ref = O.vq_forward(zr, wr, cfg, training=not eval_det, eval_deterministic=eval_det, return_scores=True)
    m = make_vq(cfg, w_cpu, eval_det)
    zg = z_cpu.to(DEV).requires_grad_(True)
    out = m(zg)
    assert out["bottleneck_rep"].dtype == torch.int64
    assert out["bottleneck_rep"].shape == ref["bottleneck_rep"].shape
    nbad = assert_indices_match(out["bottleneck_rep"], ref["bottleneck_rep"], ref["scores"], cfg.stochastic)
    if nbad:  # push the kernel's choice through the reference arithmetic for the tensor checks
        zr = z_cpu.clone().requires_grad_(True)
        wr = w_cpu.clone().requires_grad_(True)
        ref = O.vq_forward(zr, wr, cfg, training=not eval_det, eval_deterministic=eval_det,
                           forced_indices=out["bottleneck_rep"].cpu())
    for k in ("unregularized_z", "regularized_z", "emb", "loss_q", "loss_commit", "loss_codebook"):
        assert out[k].dtype == torch.float32 and out[k].shape == ref[k].shape, k
        assert_close(out[k].detach().cpu(), ref[k].detach(), k, tol)
    for k in ("loss_entropy", "per_sample_entropy", "codebook_entropy"):
        assert float(out[k]) == 0.0
    if G is None:
        G = torch.randn(z_cpu.shape, generator=torch.Generator().manual_seed(3))
    ((ref["regularized_z"] * G).sum() + g_lossq * ref["loss_q"]).backward()
    ((out["regularized_z"] * G.to(DEV)).sum() + g_lossq * out["loss_q"]).backward()
    assert zg.grad.dtype == z_cpu.dtype
    gtol = tol if z_cpu.dtype == torch.float32 else 8e-3  # grad is rounded to the latent dtype
    assert_close(zg.grad.float().cpu(), zr.grad.float(), "grad_z", gtol)
    assert_close(m.embedding.weight.grad.cpu(), wr.grad, "grad_weight", tol)
    return nbad


@pytest.mark.parametrize("name", vq_golden_names())
def test_vq_against_reference_golden(name):
    g = load_golden(name)
    cfg, eval_det = cfg_from_golden(g)
    z = golden_z(g)
    m = make_vq(cfg, g["weight"], eval_det)
    zg = z.to(DEV).requires_grad_(True)
    out = m(zg)
    # reference ranking matrix from the oracle (bit-identical to the reference on CPU)
    ref = O.vq_forward(z, torch.from_numpy(g["weight"]), cfg, training=not eval_det, eval_deterministic=eval_det,
                       return_scores=True)
    assert np.array_equal(ref["bottleneck_rep"].numpy(), g["idx"])
    nbad = assert_indices_match(out["bottleneck_rep"], g["idx"], ref["scores"], cfg.stochastic, name)
    if nbad == 0:
        for k in ("unregularized_z", "regularized_z", "emb", "loss_q", "loss_commit", "loss_codebook"):
            assert_close(out[k].detach().cpu(), g[k], k)
        G = torch.from_numpy(g["G"]).to(DEV)
        ((out["regularized_z"] * G).sum() + float(g["g_lossq"]) * out["loss_q"]).backward()
        assert_close(zg.grad.float().cpu(), g["grad_z"], "grad_z", 8e-3 if bool(g["z_is_bf16"]) else 1e-3)
        assert_close(m.embedding.weight.grad.cpu(), g["grad_weight"], "grad_weight")
    else:  # ties (edge fixtures): compare through the oracle with the kernel's own choice
        run_and_compare(z, torch.from_numpy(g["weight"]), cfg, eval_det)


@pytest.mark.parametrize("tag", ["argmin", "evaldet"])
def test_config1_indices_bit_exact(tag):
    """BASELINE configs[0] at full size (N=4096, K=8192, D=24)."""
    g = load_golden(f"config1_{tag}")
    cfg, eval_det = cfg_from_golden(g)
    m = make_vq(cfg, g["weight"], eval_det)
    out = m(torch.from_numpy(g["z"]).to(DEV))
    idx = out["bottleneck_rep"].cpu()
    ref_idx = torch.from_numpy(g["idx"].astype(np.int64))
    ref = O.vq_forward(torch.from_numpy(g["z"]), torch.from_numpy(g["weight"]), cfg, training=not eval_det,
                       eval_deterministic=eval_det, return_scores=True)
    assert_indices_match(idx, ref_idx, ref["scores"], cfg.stochastic)
    assert torch.equal(idx, ref_idx), "config 1 has no sub-1e-6 gaps (min gap 2e-5): must be bit-identical"
    assert float(out["loss_q"]) == pytest.approx(float(g["loss_q"]), rel=1e-4)


@pytest.mark.parametrize("D,K,N,dtype", [(16, 8192, 4096, torch.float32), (24, 8192, 4096, torch.float32),
                                         (16, 1000, 777, torch.bfloat16), (32, 513, 300, torch.float16),
                                         (64, 2048, 1024, torch.float32), (8, 100, 129, torch.float32),
                                         (12, 333, 200, torch.float32), (40, 257, 150, torch.bfloat16)])
@pytest.mark.parametrize("mode", ["argmin", "evaldet", "nol2"])
def test_vq_random_against_oracle(D, K, N, dtype, mode):
    torch.manual_seed(D * 1000 + K)
    cfg = O.VQConfig(dim=D, codebook_size=K, l2_normalized=mode != "nol2", stochastic=mode == "evaldet",
                     stochastic_temperature=0.03)
    z = torch.randn(1, N, D).to(dtype)
    w = torch.empty(K, D)
    torch.nn.init.kaiming_uniform_(w)
    run_and_compare(z, w, cfg, eval_det=mode == "evaldet")


def test_strided_latents_and_batch_shape():
    torch.manual_seed(4)
    cfg = O.VQConfig(dim=16, codebook_size=512, l2_normalized=True)
    big = torch.randn(3, 50, 40)
    z = big[:, 10:, 8:24]  # row stride 40, batch stride 2000 != 40*40 -> falls back to a copy
    w = torch.randn(512, 16)
    m = make_vq(cfg, w)
    out = m(z.to(DEV))
    ref = O.vq_forward(z, w, cfg, return_scores=True)
    assert_indices_match(out["bottleneck_rep"], ref["bottleneck_rep"], ref["scores"], False)
    zz = torch.randn(2, 64, 48)[:, :, :16]  # uniform row stride 48, no copy
    out = m(zz.to(DEV)[:, :, :])
    ref = O.vq_forward(zz, w, cfg, return_scores=True)
    assert_indices_match(out["bottleneck_rep"], ref["bottleneck_rep"], ref["scores"], False)
    dz = torch.randn(2, 64, 48, device=DEV)
    out = m(dz[:, :, :16])
    ref = O.vq_forward(dz[:, :, :16].cpu(), w, cfg, return_scores=True)
    assert_indices_match(out["bottleneck_rep"], ref["bottleneck_rep"], ref["scores"], False)
    assert_close(out["regularized_z"].cpu(), ref["regularized_z"], "ste")


@pytest.mark.parametrize("name", ["bottleneck_small", "bottleneck_strided_evaldet"])
def test_bottleneck_against_reference_golden(name):
    g = load_golden(name)
    B, n, C, D, K, stoch, eval_det = [int(v) for v in g["cfg"]]
    spec = {"name": "bottleneck", "args": {"bottleneck_dim": D, "norm": "none", "regularizer": {"name": "vq", "args": {
        "codebook_size": K, "commitment_loss_weight": 0.25, "codebook_loss_weight": 1.0, "entropy_loss_weight": 0.0,
        "l2_normalized": True, "stochastic": bool(stoch), "stochastic_temperature": 0.03}}}}
    m = V.make(spec, args={"token_nums": n, "input_dim": C, "output_dim": C}).to(DEV)
    assert sorted(m.state_dict().keys()) == list(g["state_keys"])
    m.load_state_dict({"in_linear.weight": torch.from_numpy(g["w_in"]), "in_linear.bias": torch.from_numpy(g["b_in"]),
                       "out_linear.weight": torch.from_numpy(g["w_out"]), "out_linear.bias": torch.from_numpy(g["b_out"]),
                       "regularizer.embedding.weight": torch.from_numpy(g["weight"])}, strict=True)
    if eval_det:
        m.eval()
        m.regularizer.set_eval_deterministic(True)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        x = torch.from_numpy(g["x"]).to(DEV).requires_grad_(True)
        out = m(x)
        assert sorted(out.keys()) == list(g["keys"])
        # indices: judged on identical latents -> feed OUR projected_z to the oracle quantizer
        cfg = O.VQConfig(dim=D, codebook_size=K, l2_normalized=True, stochastic=bool(stoch), stochastic_temperature=0.03)
        ref = O.vq_forward(out["projected_z"].detach().cpu(), torch.from_numpy(g["weight"]), cfg,
                           training=not eval_det, eval_deterministic=bool(eval_det), return_scores=True)
        assert_indices_match(out["bottleneck_rep"], ref["bottleneck_rep"], ref["scores"], bool(stoch))
        assert torch.equal(out["bottleneck_rep"].cpu(), torch.from_numpy(g["idx"]))
        for k in ("output", "projected_z", "regularized_z", "unregularized_z", "loss_q", "loss_commit"):
            assert_close(out[k].detach().cpu(), g[k], k)
        assert float(out["input_norm_first"]) == pytest.approx(float(g["input_norm_first"]), rel=1e-4)
        assert float(out["input_norm_last"]) == pytest.approx(float(g["input_norm_last"]), rel=1e-4)
        ((out["output"] * torch.from_numpy(g["Gout"]).to(DEV)).sum() + 0.1 * out["loss_q"]).backward()
        assert_close(x.grad.cpu(), g["grad_x"], "grad_x")
        assert_close(m.in_linear.weight.grad.cpu(), g["grad_w_in"], "grad_w_in")
        assert_close(m.in_linear.bias.grad.cpu(), g["grad_b_in"], "grad_b_in")
        assert_close(m.out_linear.weight.grad.cpu(), g["grad_w_out"], "grad_w_out")
        assert_close(m.out_linear.bias.grad.cpu(), g["grad_b_out"], "grad_b_out")
        assert_close(m.regularizer.embedding.weight.grad.cpu(), g["grad_weight"], "grad_weight")
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def test_decode_matches_oracle():
    torch.manual_seed(8)
    w = torch.randn(300, 24)
    w[5] = 0.0
    idx = torch.randint(0, 300, (3, 17))
    idx[0, 0] = 5
    for l2 in (True, False):
        m = V.SimpleVectorQuantizer(dim=24, codebook_size=300, l2_normalized=l2).to(DEV)
        with torch.no_grad():
            m.embedding.weight.copy_(w)
        got = m.decode(idx.to(DEV))
        ref = O.codebook_lookup(idx, w, l2)
        assert got.shape == ref.shape == (3, 17, 24)
        assert_close(got.cpu(), ref, "decode", 1e-6)
        assert torch.isfinite(got).all()
        got2 = m.get_codebook_entry(idx.to(DEV), shape=(3 * 17, 24))
        assert got2.shape == (51, 24)


@pytest.mark.parametrize("B,n,C,D,K,l2", [(3, 64, 768, 24, 1000, True), (2, 37, 512, 16, 300, False), (1, 1, 64, 8, 7, True),
                                          (5, 1024, 768, 24, 8192, True)])
def test_bottleneck_decode_fused_with_positional_table(B, n, C, D, K, l2):
    """Bottleneck.decode (bottleneck.py:166-168) + the tokenizer's `z + decoder_latent_pe`
    (larp_tokenizer.py:463-464) in one kernel: b200vq_decode_project against the oracle"""
    torch.manual_seed(B + n + K)
    spec = {"name": "bottleneck", "args": {"bottleneck_dim": D, "norm": "none", "regularizer": {"name": "vq", "args": {
        "codebook_size": K, "l2_normalized": l2, "stochastic": False}}}}
    m = V.make(spec, args={"token_nums": n, "input_dim": C, "output_dim": C}).to(DEV).eval()
    with torch.no_grad():
        m.regularizer.embedding.weight[K // 2] = 0.0          # zero-norm code row: 1e-12 clamp, never NaN
    idx = torch.randint(0, K, (B, n))
    idx[0, 0] = K // 2
    pe = torch.randn(1, n, C)
    w, wo, bo = (t.detach().cpu() for t in (m.regularizer.embedding.weight, m.out_linear.weight, m.out_linear.bias))
    ref = O.bottleneck_decode(idx, w.double(), l2, wo.double(), bo.double())
    with torch.no_grad():
        got = m.decode(idx.to(DEV))
        got_pe = m.decode(idx.to(DEV), add=pe.to(DEV))
    assert got.shape == (B, n, C) and got.dtype == torch.float32
    assert_close(got.cpu().double(), ref, "decode_project", 1e-3)
    assert_close(got_pe.cpu().double(), ref + pe.double(), "decode_project+pe", 1e-3)
    assert torch.isfinite(got).all()
    # grad-enabled call with trainable weights keeps the piecewise path and the same numbers
    got2 = m.decode(idx.to(DEV), add=pe.to(DEV))
    assert_close(got2.detach().cpu().double(), ref + pe.double(), "decode piecewise", 1e-3)


def test_decode_project_flags_bad_indices():
    from video_tokenizer_b200 import _lib
    lib = V.load_library()
    w = torch.randn(50, 16, device=DEV); wo = torch.randn(64, 16, device=DEV)
    idx = torch.tensor([3, 50, -1, 7], device=DEV)
    out = torch.zeros(4, 64, device=DEV); flag = torch.zeros(1, dtype=torch.int32, device=DEV)
    _lib.check(lib.b200vq_decode_project(idx.data_ptr(), 4, w.data_ptr(), 50, 16, 1, wo.data_ptr(), None, 64, None, 0,
                                         out.data_ptr(), 64, flag.data_ptr(), torch.cuda.current_stream().cuda_stream), "dp")
    assert int(flag) == 1
    assert torch.isnan(out[1]).all() and torch.isnan(out[2]).all() and torch.isfinite(out[0]).all() and torch.isfinite(out[3]).all()


def test_usage_stats_match_reference_golden():
    g = load_golden("code_usage_stats")
    K = int(g["K"])
    rep = torch.from_numpy(g["rep"])
    hf = torch.bincount(rep[0].flatten(), minlength=K).int().to(DEV)
    hb = torch.bincount(rep.flatten(), minlength=K).int().to(DEV)
    st = V.usage_stats(hf, hb).cpu()
    for i, k in enumerate(("index_usage", "index_usage_batch", "kl_uni", "perplexity")):
        assert float(st[i]) == pytest.approx(float(g[k]), rel=1e-4), k


def test_fused_histograms_and_usage_keys():
    torch.manual_seed(9)
    K, D, B, n = 512, 16, 3, 128
    m = V.SimpleVectorQuantizer(dim=D, codebook_size=K, l2_normalized=True, return_usage=True).to(DEV)
    out = m(torch.randn(B, n, D, device=DEV))
    rep = out["bottleneck_rep"].cpu()
    hf, hb = m.last_usage
    assert torch.equal(hb.cpu().long(), torch.bincount(rep.flatten(), minlength=K))
    assert torch.equal(hf.cpu().long(), torch.bincount(rep[0].flatten(), minlength=K))
    s = O.code_usage_stats(rep, K)
    for k in ("index_usage", "index_usage_batch", "kl_uni", "perplexity"):
        assert float(out[k]) == pytest.approx(s[k], rel=1e-4), k
    assert torch.allclose(m.code_usage().cpu(), torch.tensor([s["index_usage"], s["index_usage_batch"], s["kl_uni"],
                                                              s["perplexity"]]), rtol=1e-4)


def test_entropy_loss_variant_against_oracle():
    torch.manual_seed(10)
    cfg = O.VQConfig(dim=16, codebook_size=128, l2_normalized=True, entropy_loss_weight=0.1, entropy_loss_temperature=0.01)
    z = torch.randn(2, 32, 16)
    w = torch.randn(128, 16)
    zr, wr = z.clone().requires_grad_(True), w.clone().requires_grad_(True)
    ref = O.vq_forward(zr, wr, cfg)
    m = V.SimpleVectorQuantizer(dim=16, codebook_size=128, l2_normalized=True, entropy_loss_weight=0.1).to(DEV)
    with torch.no_grad():
        m.embedding.weight.copy_(w)
    zg = z.to(DEV).requires_grad_(True)
    out = m(zg)
    assert torch.equal(out["bottleneck_rep"].cpu(), ref["bottleneck_rep"])
    for k in ("loss_q", "loss_entropy", "per_sample_entropy", "codebook_entropy"):
        assert_close(out[k].detach().cpu(), ref[k].detach(), k)
    ref["loss_q"].backward()
    out["loss_q"].backward()
    assert_close(zg.grad.cpu(), zr.grad, "grad_z (entropy)", 2e-3)
    assert_close(m.embedding.weight.grad.cpu(), wr.grad, "grad_w (entropy)", 2e-3)
    ms = V.SimpleVectorQuantizer(dim=16, codebook_size=128, l2_normalized=True, stochastic=True, entropy_loss_weight=0.1).to(DEV)
    with pytest.raises(UnboundLocalError):
        ms(z.to(DEV))


def test_errors_are_loud():
    m = V.SimpleVectorQuantizer(dim=16, codebook_size=64, l2_normalized=True)
    with pytest.raises(RuntimeError):
        m(torch.randn(1, 4, 16))  # CPU tensors: no fallback
    m = m.to(DEV)
    with pytest.raises(AssertionError):
        m(torch.randn(4, 16, device=DEV))
    with pytest.raises(ValueError):
        m(torch.randn(1, 4, 24, device=DEV))
    with pytest.raises(AssertionError):
        V.SimpleVectorQuantizer(dim=16, codebook_size=64, l2_normalized=False, stochastic=True).to(DEV)(
            torch.randn(1, 4, 16, device=DEV))


def test_decode_is_differentiable_towards_the_codebook_when_grad_is_recorded():
    """reference get_codebook_entry (bottleneck.py:327-341) is nn.Embedding + F.normalize: gradients reach the
    codebook.  The kernel path builds no graph, so with grad enabled the drop-in uses the same op sequence."""
    torch.manual_seed(31)
    m = V.SimpleVectorQuantizer(dim=16, codebook_size=64, l2_normalized=True).to(DEV)
    idx = torch.randint(0, 64, (2, 5), device=DEV)
    with torch.no_grad():
        fast = m.decode(idx)
    slow = m.decode(idx)
    assert slow.requires_grad and tuple(slow.shape) == (2, 5, 16)
    assert_close(slow.detach(), fast, "decode grad path vs kernel path", tol=1e-6)
    slow.square().sum().backward()
    ref_w = m.embedding.weight.detach().clone().requires_grad_(True)
    F.normalize(F.embedding(idx, ref_w), dim=-1).square().sum().backward()
    assert torch.allclose(m.embedding.weight.grad, ref_w.grad, atol=1e-6)
    assert tuple(m.get_codebook_entry(idx, shape=(10, 16)).shape) == (10, 16)

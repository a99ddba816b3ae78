"""Entropy regulariser of the quantizer (reference models/bottleneck.py:12-33 as called at :300-304; SURVEY.md 8a-8 /
8f N4): the oracle against golden vectors generated from the unmodified reference class
(GOLDEN_ONLY=entropy python -m oracle.make_golden), the closed-form streaming formulation against autograd, and
the sm_100a streaming kernels (b200vq_entropy_forward / _backward) against both."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import vq_oracle as O
from tests.helpers import GOLDEN, assert_close, load_golden

ENT_GOLDEN = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "ent_*.npz")))
KEYS = ("loss_q", "loss_commit", "loss_codebook", "loss_entropy", "per_sample_entropy", "codebook_entropy")


def _cfg(g):
    D, K, l2 = (int(v) for v in g["cfg"])
    w_ent, T_ent = (float(v) for v in g["cfg_f"])
    return O.VQConfig(dim=D, codebook_size=K, l2_normalized=bool(l2), entropy_loss_weight=w_ent,
                      entropy_loss_temperature=T_ent)


def _objective(out, G):
    return ((out["regularized_z"] * G).sum() + 0.7 * out["loss_q"] + 0.3 * out["per_sample_entropy"]
            - 0.2 * out["codebook_entropy"])


def test_entropy_golden_present():
    assert len(ENT_GOLDEN) == 4


@pytest.mark.parametrize("name", ENT_GOLDEN)
def test_entropy_oracle_matches_reference_golden(name):
    g = load_golden(name)
    cfg = _cfg(g)
    z = torch.from_numpy(g["z"]).requires_grad_(True)
    w = torch.from_numpy(g["weight"]).clone().requires_grad_(True)
    out = O.vq_forward(z, w, cfg)
    assert np.array_equal(out["bottleneck_rep"].numpy(), g["idx"])
    for k in KEYS:
        assert float(out[k].detach()) == pytest.approx(float(g[k]), rel=1e-6, abs=1e-9), k
    _objective(out, torch.from_numpy(g["G"])).backward()
    assert_close(z.grad, g["grad_z"], "grad_z", 1e-6)
    assert_close(w.grad, g["grad_weight"], "grad_weight", 1e-6)


@pytest.mark.parametrize("name", ENT_GOLDEN)
def test_entropy_closed_form_matches_autograd(name):
    """the streaming formulation (float64) against autograd of the reference expression on the same tensors"""
    g = load_golden(name)
    cfg = _cfg(g)
    z = torch.from_numpy(g["z"]).reshape(-1, cfg.dim)
    w = torch.from_numpy(g["weight"])
    if cfg.l2_normalized:
        z = torch.nn.functional.normalize(z, dim=-1)
        w = torch.nn.functional.normalize(w, dim=-1)
    zr, er = z.clone().requires_grad_(True), w.clone().requires_grad_(True)
    d = (torch.sum(zr ** 2, dim=1, keepdim=True) + torch.sum(er ** 2, dim=1) - 2 * zr @ er.t())
    loss, sample, avg = O.entropy_terms(-d, temperature=cfg.entropy_loss_temperature)
    (0.9 * loss + 0.3 * sample - 0.2 * avg).backward()
    (l2, s2, a2), gz, ge = O.entropy_closed_form(z, w, cfg.entropy_loss_temperature, 0.9, 0.3, -0.2)
    assert float(l2) == pytest.approx(float(loss), rel=2e-5, abs=1e-6)
    assert float(s2) == pytest.approx(float(sample), rel=2e-5, abs=1e-6)
    assert float(a2) == pytest.approx(float(avg), rel=2e-5, abs=1e-6)
    assert_close(gz.float(), zr.grad, "closed-form grad_z", 1e-4)
    assert_close(ge.float(), er.grad, "closed-form grad_emb", 1e-4)


# ---------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("name", ENT_GOLDEN)
def test_entropy_dropin_matches_reference_golden(name):
    import video_tokenizer_b200 as V
    from video_tokenizer_b200 import functional as VF
    g = load_golden(name)
    cfg = _cfg(g)
    assert VF.entropy_supported(cfg.dim)
    m = V.SimpleVectorQuantizer(dim=cfg.dim, codebook_size=cfg.codebook_size, l2_normalized=cfg.l2_normalized,
                                entropy_loss_weight=cfg.entropy_loss_weight,
                                entropy_loss_temperature=cfg.entropy_loss_temperature).cuda()
    with torch.no_grad():
        m.embedding.weight.copy_(torch.from_numpy(g["weight"]))
    z = torch.from_numpy(g["z"]).cuda().requires_grad_(True)
    out = m(z)
    assert np.array_equal(out["bottleneck_rep"].cpu().numpy(), g["idx"])
    for k in KEYS:
        assert float(out[k].detach()) == pytest.approx(float(g[k]), rel=1e-3, abs=1e-6), k
    _objective(out, torch.from_numpy(g["G"]).cuda()).backward()
    assert_close(z.grad.cpu(), g["grad_z"], "grad_z", 1e-3)
    assert_close(m.embedding.weight.grad.cpu(), g["grad_weight"], "grad_weight", 1e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("N,K,D,T", [(1, 8, 8, 0.5), (127, 65, 16, 0.05), (300, 2049, 24, 0.01), (4096, 8192, 24, 0.01),
                                     (513, 129, 32, 0.1), (40000, 300, 16, 0.05), (200, 40000, 8, 0.05),
                                     (300, 513, 64, 0.05)])
def test_entropy_kernels_match_closed_form(N, K, D, T):
    """the kernels alone (ragged N and K, single row, the config-1 size; split code streams, one and two owner
    vectors per thread on either side) against the float64 closed form, with separate upstream gradients on the
    three outputs"""
    from video_tokenizer_b200 import functional as VF
    gen = torch.Generator().manual_seed(N * 7 + K)
    z = torch.nn.functional.normalize(torch.randn(N, D, generator=gen), dim=-1)
    e = torch.nn.functional.normalize(torch.randn(K, D, generator=gen), dim=-1)
    zg, eg = z.cuda().requires_grad_(True), e.cuda().requires_grad_(True)
    loss, sample, avg = VF.entropy_loss_streamed(zg, eg, T)
    (0.9 * loss + 0.3 * sample - 0.2 * avg).backward()
    (l2, s2, a2), gz, ge = O.entropy_closed_form(z, e, T, 0.9, 0.3, -0.2)
    assert float(loss) == pytest.approx(float(l2), rel=1e-3, abs=1e-5)
    assert float(sample) == pytest.approx(float(s2), rel=1e-3, abs=1e-5)
    assert float(avg) == pytest.approx(float(a2), rel=1e-3, abs=1e-5)
    assert_close(zg.grad.cpu(), gz.float(), "grad_z", 1e-3)
    assert_close(eg.grad.cpu(), ge.float(), "grad_emb", 1e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("N,K,D", [(2000, 1000, 24), (65536, 8192, 24)])
def test_entropy_streamed_equals_torch_formulation_on_device(N, K, D):
    """same device, same tensors: the streaming kernels against the reference's N x K op sequence in torch
    (the second case is BASELINE's full bench size: 64 x 1024 tokens, K = 8192 -- 2 GiB per matrix for torch)"""
    import video_tokenizer_b200 as V
    from video_tokenizer_b200 import functional as VF
    torch.manual_seed(5)
    z = torch.nn.functional.normalize(torch.randn(N, D), dim=-1).cuda()
    e = torch.nn.functional.normalize(torch.randn(K, D), dim=-1).cuda()
    z1, e1 = z.clone().requires_grad_(True), e.clone().requires_grad_(True)
    z2, e2 = z.clone().requires_grad_(True), e.clone().requires_grad_(True)
    a = VF.entropy_loss_streamed(z1, e1, 0.01)
    d = torch.sum(z2 ** 2, dim=1, keepdim=True) + torch.sum(e2 ** 2, dim=1) - 2 * z2 @ e2.t()
    b = V.entropy_loss(-d, temperature=0.01)
    for x, y in zip(a, b):
        assert float(x) == pytest.approx(float(y), rel=1e-3, abs=1e-5)
    a[0].backward()
    b[0].backward()
    assert_close(z1.grad, z2.grad, "grad_z", 1e-3)
    assert_close(e1.grad, e2.grad, "grad_emb", 1e-3)


@pytest.mark.gpu
def test_entropy_unsupported_dim_uses_the_matrix_formulation():
    import video_tokenizer_b200 as V
    from video_tokenizer_b200 import functional as VF
    assert VF.entropy_supported(64) and not VF.entropy_supported(12) and not VF.entropy_supported(128)
    m = V.SimpleVectorQuantizer(dim=12, codebook_size=50, l2_normalized=True, entropy_loss_weight=0.1).cuda()
    out = m(torch.randn(1, 10, 12, device="cuda"))
    assert torch.isfinite(out["loss_entropy"])

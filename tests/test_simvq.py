"""Sibling quantizer `SimVQ` (reference models/cosmos.py:772-876): the oracle restatement against
golden vectors generated from the unmodified reference class (GOLDEN_ONLY=simvq python -m oracle.make_golden), and
the sm_100a drop-in against both."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import vq_oracle as O
from tests.helpers import GOLDEN, assert_close, load_golden

SIMVQ_GOLDEN = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "simvq_*.npz")))


def _t(g, k, grad=False):
    return torch.from_numpy(g[k]).clone().requires_grad_(grad)


def test_simvq_golden_present():
    assert len(SIMVQ_GOLDEN) == 4


@pytest.mark.parametrize("name", SIMVQ_GOLDEN)
def test_simvq_oracle_matches_reference_golden(name):
    g = load_golden(name)
    z, pw, pb = _t(g, "z", True), _t(g, "proj_w", True), _t(g, "proj_b", True)
    out = O.simvq_forward(z, _t(g, "anchors"), pw, pb, float(g["beta"]), bool(g["legacy"]))
    assert np.array_equal(out["indices"].numpy(), g["indices"])
    assert np.array_equal(out["z_q"].detach().numpy(), g["z_q"])               # same torch ops: bit-exact
    assert float(out["loss"].detach()) == pytest.approx(float(g["loss"]), rel=1e-6)
    ((out["z_q"] * _t(g, "G")).sum() + 0.7 * out["loss"]).backward()
    assert_close(z.grad, g["grad_z"], "grad_z", 1e-6)
    assert_close(pw.grad, g["grad_proj_w"], "grad_proj_w", 1e-6)
    assert_close(pb.grad, g["grad_proj_b"], "grad_proj_b", 1e-6)


def _dropin(g):
    import video_tokenizer_b200 as V
    K, D = g["anchors"].shape
    q = V.SimVQ(n_e=K, e_dim=D, beta=float(g["beta"]), legacy=bool(g["legacy"]))
    with torch.no_grad():
        q.embedding.weight.copy_(_t(g, "anchors"))
        q.embedding_proj.weight.copy_(_t(g, "proj_w"))
        q.embedding_proj.bias.copy_(_t(g, "proj_b"))
    return q.cuda().train()


@pytest.mark.gpu
@pytest.mark.parametrize("name", SIMVQ_GOLDEN)
def test_simvq_dropin_matches_reference_golden(name):
    g = load_golden(name)
    q = _dropin(g)
    assert not q.embedding.weight.requires_grad
    assert sorted(q.state_dict()) == ["embedding.weight", "embedding_proj.bias", "embedding_proj.weight"]
    z = _t(g, "z").cuda().requires_grad_(True)
    z_q, loss, (perp, ind) = q(z)
    assert perp is None and z_q.shape == z.shape and ind.dtype == torch.int64
    assert ind.shape == tuple(g["indices"].shape)
    ref = O.simvq_forward(_t(g, "z"), _t(g, "anchors"), _t(g, "proj_w"), _t(g, "proj_b"), float(g["beta"]),
                          bool(g["legacy"]))
    got = ind.reshape(-1).cpu()
    mism = (got != torch.from_numpy(g["indices"]).reshape(-1)).nonzero().flatten()
    for r in mism.tolist():     # only rows whose two best codes are closer than 1e-6 in the reference's own matrix
        d = ref["scores"][r]
        assert float(d[got[r]] - d.min()) < 1e-6
    assert len(mism) <= 2
    assert_close(z_q.detach().cpu(), g["z_q"], "z_q", 1e-3 if len(mism) else 1e-5)
    assert float(loss.detach()) == pytest.approx(float(g["loss"]), rel=1e-4)
    ((z_q * _t(g, "G").cuda()).sum() + 0.7 * loss).backward()
    assert q.embedding.weight.grad is None
    assert_close(z.grad.cpu(), g["grad_z"], "grad_z", 1e-3)
    assert_close(q.embedding_proj.weight.grad.cpu(), g["grad_proj_w"], "grad_proj_w", 1e-3)
    assert_close(q.embedding_proj.bias.grad.cpu(), g["grad_proj_b"], "grad_proj_b", 1e-3)
    # decode path
    with torch.no_grad():
        ent = q.get_codebook_entry(torch.from_numpy(g["entry_indices"]).cuda())
    assert ent.shape == tuple(g["entry"].shape)
    assert_close(ent.cpu(), g["entry"], "get_codebook_entry", 1e-5)
    assert int(q.last_usage.sum()) == got.numel()


@pytest.mark.gpu
def test_simvq_at_the_reference_tokenizer_dim():
    """e_dim = 1024 as the reference builds it (cosmos.py:918-924), against the oracle on the same tensors"""
    import video_tokenizer_b200 as V
    torch.manual_seed(3)
    q = V.SimVQ(n_e=777, e_dim=1024).cuda()
    z = (torch.randn(1, 1024, 2, 5, 5) * 0.5)
    zg = z.cuda().requires_grad_(True)
    z_q, loss, (_, ind) = q(zg)
    zr = z.clone().requires_grad_(True)
    pw = q.embedding_proj.weight.detach().cpu().requires_grad_(True)
    pb = q.embedding_proj.bias.detach().cpu().requires_grad_(True)
    ref = O.simvq_forward(zr, q.embedding.weight.detach().cpu(), pw, pb, 0.25, True)
    got = ind.reshape(-1).cpu()
    for r in (got != ref["indices"].reshape(-1)).nonzero().flatten().tolist():
        d = ref["scores"][r]
        assert float(d[got[r]] - d.min()) < 1e-4 * float(d.min().abs() + 1)     # fp32 rounding of a 1024-term sum
    assert float(loss.detach()) == pytest.approx(float(ref["loss"].detach()), rel=1e-3)
    G = torch.randn(z_q.shape)
    ((z_q * G.cuda()).sum() + loss).backward()
    ((ref["z_q"] * G).sum() + ref["loss"]).backward()
    assert_close(zg.grad.cpu(), zr.grad, "grad_z", 1e-3)
    assert_close(q.embedding_proj.weight.grad.cpu(), pw.grad, "grad_proj_w", 2e-3)


@pytest.mark.gpu
def test_simvq_rejects_bad_shapes():
    import video_tokenizer_b200 as V
    q = V.SimVQ(n_e=32, e_dim=16).cuda()
    with pytest.raises(ValueError):
        q(torch.randn(2, 8, 4, 4, device="cuda"))
    with pytest.raises(ValueError):
        q(torch.randn(2, 16, 4, device="cuda"))

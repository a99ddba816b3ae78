"""Sibling quantizer 'sq' (reference models/model_new/quantizer/fsq.py:148-230):
the oracle restatement against golden vectors generated from the unmodified reference class
(oracle/make_golden.py, GOLDEN_ONLY=sq), and the sm_100a drop-in against both."""
import numpy as np
import pytest
import torch

from oracle import vq_oracle as O
from tests.helpers import assert_close, load_golden, rel_err

SQ_GOLDEN = ["sq_l2_d24_k1000_blc", "sq_l2_d24_k4096_frozen", "sq_nol2_d16_k300_bchw"]


def _oracle_run(g):
    z = torch.from_numpy(g["z"]).requires_grad_(True)
    W = torch.from_numpy(g["weight"]).requires_grad_(True)
    out = O.sq_forward(z, W, bool(g["l2"]), float(g["beta"]), str(g["input_format"]))
    ((out["output"] * torch.from_numpy(g["G"])).sum() + 0.7 * out["loss_codebook"]).backward()
    return z, W, out


@pytest.mark.parametrize("name", SQ_GOLDEN)
def test_sq_oracle_matches_reference_golden(name):
    g = load_golden(name)
    z, W, out = _oracle_run(g)
    assert np.array_equal(out["output"].detach().numpy(), g["output"])          # same torch ops: bit-exact
    assert float(out["loss_codebook"]) == pytest.approx(float(g["loss_codebook"]), rel=1e-6)
    assert_close(z.grad, g["grad_z"], "grad_z", 1e-6)
    if not bool(g["frozen"]):
        assert_close(W.grad, g["grad_weight"], "grad_weight", 1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("name", SQ_GOLDEN)
def test_sq_dropin_matches_reference_golden(name, tmp_path):
    import video_tokenizer_b200 as V
    g = load_golden(name)
    K, D = g["weight"].shape
    path = str(tmp_path / "cb.npy")
    np.save(path, g["weight"])
    q = V.VectorQuantizer(n_embed=K, embed_dim=D, l2_norm=bool(g["l2"]), beta=float(g["beta"]),
                          input_format=str(g["input_format"]), predefined_codebook=path,
                          freeze_codebook=bool(g["frozen"])).cuda().train()
    assert q.embedding.weight.requires_grad == (not bool(g["frozen"]))
    z = torch.from_numpy(g["z"]).cuda().requires_grad_(True)
    out = q(z)
    assert set(out.keys()) == {"output", "loss_codebook"}
    assert out["output"].shape == z.shape
    # indices bit-exact against the oracle's ranking (ties -> lowest index), outputs inside 1e-3
    ref = O.sq_forward(torch.from_numpy(g["z"]), torch.from_numpy(g["weight"]), bool(g["l2"]), float(g["beta"]),
                       str(g["input_format"]))
    mism = (q.last_indices.cpu() != ref["indices"]).nonzero().flatten()
    for r in mism.tolist():      # only rows whose two best codes are closer than 1e-6 in the reference's own matrix
        d = ref["scores"][r]
        assert float(d[q.last_indices[r]] - d.min()) < 1e-6
    assert len(mism) <= 2
    assert_close(out["output"].detach().cpu(), g["output"], "output", 1e-3 if len(mism) else 1e-5)
    assert float(out["loss_codebook"]) == pytest.approx(float(g["loss_codebook"]), rel=1e-4)
    ((out["output"] * torch.from_numpy(g["G"]).cuda()).sum() + 0.7 * out["loss_codebook"]).backward()
    assert_close(z.grad.cpu(), g["grad_z"], "grad_z", 1e-3)
    if not bool(g["frozen"]):
        assert_close(q.embedding.weight.grad.cpu(), g["grad_weight"], "grad_weight", 1e-3)
    else:
        assert q.embedding.weight.grad is None
    assert float(q.usage_entropy()) == pytest.approx(float(ref["cb_entropy"]), rel=1e-4)
    # decode
    idx = q.last_indices.reshape(g["z"].shape[0], -1)
    W = torch.from_numpy(g["weight"])
    want = torch.nn.functional.normalize(W[idx.cpu()], dim=-1) if bool(g["l2"]) else W[idx.cpu()]
    if str(g["input_format"]) == "bchw":
        with pytest.raises(AssertionError):          # 30 positions are not a square grid (fsq.py:227-228)
            q.get_codebook_entry(idx)
        ent = q.get_codebook_entry(idx[:, :25])
        assert ent.shape == (idx.shape[0], D, 5, 5)
        assert_close(ent.permute(0, 2, 3, 1).reshape(idx.shape[0], 25, D).cpu(), want[:, :25], "get_codebook_entry", 1e-6)
    else:
        assert_close(q.get_codebook_entry(idx).cpu(), want, "get_codebook_entry", 1e-6)


@pytest.mark.gpu
def test_sq_full_size_table_properties():
    """the 'sq' configuration itself: K = 196 560 unit codes, D = 24 (larp_tokenizer.py:233); the reference
    would build a 196 560-wide row per token.  Size-independent checks: chunked fp32 brute force on a
    row sample, idempotence (a code row quantizes to itself), loss identity for unit vectors."""
    import video_tokenizer_b200 as V
    K, D, N = 196_560, 24, 16384
    torch.manual_seed(5)
    W = torch.nn.functional.normalize(torch.randn(K, D, device="cuda"), dim=-1)
    q = V.VectorQuantizer(n_embed=K, embed_dim=D, l2_norm=True, beta=0.25, input_format="blc", predefined_codebook=None).cuda()
    with torch.no_grad():
        q.embedding.weight.copy_(W)
    q.embedding.weight.requires_grad = False
    z = torch.randn(16, N // 16, D, device="cuda", requires_grad=True)
    out = q(z)
    idx = q.last_indices
    assert idx.shape == (N,) and int(idx.min()) >= 0 and int(idx.max()) < K
    zn = torch.nn.functional.normalize(z.detach().reshape(-1, D), dim=-1)
    rows = torch.arange(0, N, 37, device="cuda")
    s = zn[rows] @ W.t()                                         # (443, K) fp32
    best = s.max(dim=1)
    picked = s.gather(1, idx[rows, None]).squeeze(1)
    assert float((best.values - picked).max()) < 1e-6           # the picked code is a maximiser ...
    exact = best.indices == idx[rows]
    assert float(exact.float().mean()) > 0.99                   # ... and the same index except at fp32 near-ties
    # loss = (1 + beta) * mean_i (2 - 2 s_i) for unit vectors
    s_sel = (zn * W[idx]).sum(-1)
    assert float(out["loss_codebook"]) == pytest.approx(float(1.25 * (2 - 2 * s_sel).mean()), rel=1e-4)
    # idempotence: code rows quantize to themselves
    probe = torch.randint(0, K, (4096,), device="cuda")
    out2 = q(W[probe].reshape(4, 1024, D))
    assert torch.equal(q.last_indices, probe) or float((q.last_indices == probe).float().mean()) > 0.999
    assert rel_err(out2["output"].reshape(-1, D), W[probe]) < 1e-6
    out["loss_codebook"].backward()
    assert z.grad is not None and torch.isfinite(z.grad).all()

"""CPU: the oracle restatement against every golden vector produced by the reference."""
import numpy as np
import pytest
import torch

from oracle import vq_oracle as O
from oracle.ref_loader import reference_root
from tests.helpers import (assert_close, cfg_from_golden, golden_z, load_golden, vq_golden_names)


@pytest.mark.parametrize("name", vq_golden_names())
def test_vq_oracle_matches_reference_golden(name):
    g = load_golden(name)
    cfg, eval_det = cfg_from_golden(g)
    z = golden_z(g).clone().requires_grad_(True)
    w = torch.from_numpy(g["weight"]).clone().requires_grad_(True)
    out = O.vq_forward(z, w, cfg, training=not eval_det, eval_deterministic=eval_det, return_scores=True)
    assert np.array_equal(out["bottleneck_rep"].numpy(), g["idx"]), "oracle indices differ from the reference's"
    assert out["bottleneck_rep"].dtype == torch.int64
    for k in ("loss_q", "loss_commit", "loss_codebook"):
        assert float(out[k]) == pytest.approx(float(g[k]), rel=1e-6, abs=1e-9)
    for k in ("unregularized_z", "regularized_z", "emb"):
        assert np.array_equal(out[k].detach().numpy(), g[k]), k
    G = torch.from_numpy(g["G"])
    ((out["regularized_z"] * G).sum() + float(g["g_lossq"]) * out["loss_q"]).backward()
    assert_close(z.grad.float(), g["grad_z"], "grad_z", 1e-6)
    assert_close(w.grad, g["grad_weight"], "grad_weight", 1e-6)
    assert np.allclose(out["top2_gap"].numpy(), g["top2_gap"], rtol=0, atol=0)
    # closed-form backward (float64) referees the autograd result
    dz, dW = O.closed_form_backward(golden_z(g).float(), torch.from_numpy(g["weight"]), out["bottleneck_rep"],
                                    cfg, G, float(g["g_lossq"]))
    assert_close(dz, g["grad_z"], "closed-form grad_z", 2e-3 if bool(g["z_is_bf16"]) else 1e-5)
    assert_close(dW, g["grad_weight"], "closed-form grad_weight", 1e-5)


@pytest.mark.parametrize("name", ["bottleneck_small", "bottleneck_strided_evaldet"])
def test_bottleneck_oracle_matches_reference_golden(name):
    g = load_golden(name)
    B, n, C, D, K, stoch, eval_det = [int(v) for v in g["cfg"]]
    cfg = O.VQConfig(dim=D, codebook_size=K, l2_normalized=True, stochastic=bool(stoch), stochastic_temperature=0.03)
    t = {k: torch.from_numpy(g[k]).clone().requires_grad_(True)
         for k in ("x", "w_in", "b_in", "w_out", "b_out", "weight")}
    out = O.bottleneck_forward(t["x"], t["w_in"], t["b_in"], t["w_out"], t["b_out"], t["weight"], cfg,
                               training=not eval_det, eval_deterministic=bool(eval_det))
    assert sorted(out.keys()) == list(g["keys"])
    assert np.array_equal(out["bottleneck_rep"].numpy(), g["idx"])
    for k in ("output", "projected_z", "regularized_z", "unregularized_z"):
        assert np.array_equal(out[k].detach().numpy(), g[k]), k
    assert out["input_norm_first"] == pytest.approx(float(g["input_norm_first"]), rel=1e-6)
    assert out["input_norm_last"] == pytest.approx(float(g["input_norm_last"]), rel=1e-6)
    ((out["output"] * torch.from_numpy(g["Gout"])).sum() + 0.1 * out["loss_q"]).backward()
    for k, gk in (("x", "grad_x"), ("w_in", "grad_w_in"), ("b_in", "grad_b_in"), ("w_out", "grad_w_out"),
                  ("b_out", "grad_b_out"), ("weight", "grad_weight")):
        assert_close(t[k].grad, g[gk], gk, 1e-6)


@pytest.mark.parametrize("tag", ["argmin", "evaldet"])
def test_config1_oracle_matches_reference_golden(tag):
    """BASELINE configs[0]: B=4x1024, D=24, K=8192 (cfgs/larp_tokenizer.yaml block)."""
    g = load_golden(f"config1_{tag}")
    cfg, eval_det = cfg_from_golden(g)
    out = O.vq_forward(torch.from_numpy(g["z"]), torch.from_numpy(g["weight"]), cfg, training=not eval_det,
                       eval_deterministic=eval_det, return_scores=True)
    assert np.array_equal(out["bottleneck_rep"].numpy().astype(np.int16), g["idx"])
    assert float(out["loss_q"]) == pytest.approx(float(g["loss_q"]), rel=1e-6)
    assert np.array_equal(out["top2_gap"].numpy(), g["top2_gap"])


@pytest.mark.parametrize("name", ["stoch_probs_d16_k512", "stoch_probs_d24_k8192"])
def test_stochastic_probabilities_match_reference(name):
    g = load_golden(name)
    D, K = g["weight"].shape[1], g["weight"].shape[0]
    cfg = O.VQConfig(dim=D, codebook_size=K, l2_normalized=True, stochastic=True, stochastic_temperature=0.03)
    out = O.vq_forward(torch.from_numpy(g["z"]), torch.from_numpy(g["weight"]), cfg, return_scores=True,
                       generator=torch.Generator().manual_seed(0))
    assert np.array_equal(out["scores"].numpy(), g["cos"])
    probs = torch.softmax(out["scores"] * float(g["t_inv"]), dim=-1)
    assert np.array_equal(probs.numpy(), g["probs"])
    # the oracle's multinomial draws follow those probabilities (sanity of the sampling branch)
    assert int(out["bottleneck_rep"].max()) < K


def test_code_usage_stats_match_reference():
    g = load_golden("code_usage_stats")
    s = O.code_usage_stats(torch.from_numpy(g["rep"]), int(g["K"]))
    for k in ("index_usage", "index_usage_batch", "kl_uni", "perplexity"):
        assert s[k] == pytest.approx(float(g[k]), rel=1e-5), k


def test_stochastic_entropy_error_parity():
    """Reference raises UnboundLocalError for stochastic + entropy weight > 0 (bottleneck.py:272-301)."""
    cfg = O.VQConfig(dim=8, codebook_size=16, l2_normalized=True, stochastic=True, entropy_loss_weight=0.1)
    with pytest.raises(UnboundLocalError):
        O.vq_forward(torch.randn(1, 4, 8), torch.randn(16, 8), cfg)


@pytest.mark.skipif(reference_root() is None, reason="reference tree only exists in the build container")
def test_live_reference_crosscheck():
    """Container only: oracle vs the live reference module on fresh random inputs."""
    from oracle.ref_loader import load_reference
    bott, _ = load_reference()
    for seed, (D, K, stoch) in enumerate([(16, 300, False), (24, 513, True)]):
        torch.manual_seed(100 + seed)
        vq = bott.SimpleVectorQuantizer(dim=D, codebook_size=K, l2_normalized=True, stochastic=stoch,
                                        stochastic_temperature=0.03)
        z = torch.randn(2, 70, D)
        cfg = O.VQConfig(dim=D, codebook_size=K, l2_normalized=True, stochastic=stoch, stochastic_temperature=0.03)
        torch.manual_seed(7)
        ref = vq(z)
        torch.manual_seed(7)
        mine = O.vq_forward(z, vq.embedding.weight.detach(), cfg, training=True)
        assert torch.equal(ref["bottleneck_rep"], mine["bottleneck_rep"])  # multinomial draws included
        assert torch.equal(ref["regularized_z"], mine["regularized_z"])
        assert float(ref["loss_q"]) == float(mine["loss_q"])

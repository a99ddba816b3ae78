"""GPU: stochastic (SVQ) mode is validated distributionally, as BASELINE.json's north_star
asks: chi-square of sampled-index histograms against the REFERENCE's softmax probabilities
(golden, bottleneck.py:275-276), plus the expected commitment loss under that distribution."""
import numpy as np
import pytest
import torch
from scipy import stats

import video_tokenizer_b200 as V
from tests.helpers import load_golden

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _sample(g, M, seed):
    z = torch.from_numpy(g["z"])[0]            # (R, D)
    R, D = z.shape
    K = g["weight"].shape[0]
    m = V.SimpleVectorQuantizer(dim=D, codebook_size=K, l2_normalized=True, stochastic=True,
                                stochastic_temperature=0.03).to(DEV)
    with torch.no_grad():
        m.embedding.weight.copy_(torch.from_numpy(g["weight"]))
    m.train()
    torch.cuda.manual_seed(seed)
    zrep = z.to(DEV).repeat_interleave(M, dim=0).reshape(1, R * M, D)  # noise is keyed by row -> independent draws
    out = m(zrep)
    return out, out["bottleneck_rep"].reshape(R, M).cpu(), K


@pytest.mark.parametrize("name,M", [("stoch_probs_d16_k512", 20000), ("stoch_probs_d24_k8192", 20000)])
def test_chi_square_against_reference_probabilities(name, M):
    g = load_golden(name)
    out, draws, K = _sample(g, M, seed=1234)
    probs = g["probs"].astype(np.float64)
    R = probs.shape[0]
    pvals = []
    for r in range(R):
        counts = np.bincount(draws[r].numpy(), minlength=K).astype(np.float64)
        p = probs[r] / probs[r].sum()
        order = np.argsort(-p)
        exp = p[order] * M
        obs = counts[order]
        big = exp >= 5.0
        e = np.concatenate([exp[big], [exp[~big].sum()]])
        o = np.concatenate([obs[big], [obs[~big].sum()]])
        if e[-1] < 5.0 and len(e) > 2:  # fold a too-small tail into the last kept bin
            e[-2] += e[-1]; o[-2] += o[-1]; e = e[:-1]; o = o[:-1]
        chi2 = ((o - e) ** 2 / e).sum()
        pvals.append(stats.chi2.sf(chi2, df=len(e) - 1))
    # Bonferroni over rows at 1e-3
    assert min(pvals) > 1e-3 / R, f"chi-square rejects: p-values {pvals}"
    # expected commitment loss under the reference distribution
    assert float(out["loss_commit"]) == pytest.approx(float(g["expected_loss_commit"]), rel=5e-3)


def test_draws_change_with_generator_and_repeat_with_seed():
    g = load_golden("stoch_probs_d16_k512")
    _, a, _ = _sample(g, 256, seed=7)
    _, b, _ = _sample(g, 256, seed=7)
    _, c, _ = _sample(g, 256, seed=8)
    assert torch.equal(a, b), "same CUDA generator state must reproduce the draws (checkpoint/resume contract)"
    assert not torch.equal(a, c)
    # consecutive calls advance the generator offset
    z = torch.from_numpy(g["z"]).to(DEV).repeat(1, 64, 1)
    m = V.SimpleVectorQuantizer(dim=16, codebook_size=512, l2_normalized=True, stochastic=True,
                                stochastic_temperature=0.03).to(DEV)
    r1 = m(z)["bottleneck_rep"]
    r2 = m(z)["bottleneck_rep"]
    assert not torch.equal(r1, r2)


def test_eval_deterministic_switch():
    g = load_golden("stoch_probs_d16_k512")
    z = torch.from_numpy(g["z"]).to(DEV)
    m = V.SimpleVectorQuantizer(dim=16, codebook_size=512, l2_normalized=True, stochastic=True,
                                stochastic_temperature=0.03).to(DEV)
    with torch.no_grad():
        m.embedding.weight.copy_(torch.from_numpy(g["weight"]))
    m.eval()
    m.set_eval_deterministic(True)
    idx = m(z)["bottleneck_rep"].cpu().flatten()
    assert torch.equal(idx, torch.from_numpy(g["probs"]).argmax(1))
    m.set_stochastic_temperature(0.5)
    assert m.stochastic_temperature_inv == 2.0

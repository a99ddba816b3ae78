"""Shared test helpers: golden loading and the parity rules of BASELINE.json's north_star."""
import glob
import os

import numpy as np
import torch

from oracle import vq_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REL_TOL = 1e-3   # north_star: outputs, losses, grads within 1e-3 relative error
GAP_TOL = 1e-6   # north_star: index mismatch only where the reference top-2 gap < 1e-6


def load_golden(name):
    d = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    return {k: d[k] for k in d.files}


def vq_golden_names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "vq_*.npz")))


def golden_z(g):
    z = torch.from_numpy(g["z"])
    if "z_is_bf16" in g and bool(g["z_is_bf16"]):
        z = z.view(torch.bfloat16)
    return z


def cfg_from_golden(g):
    D, K, l2, stoch, eval_det, same = [int(v) for v in g["cfg"]]
    beta, w_cb, T = [float(v) for v in g["cfg_f"]]
    cfg = O.VQConfig(dim=D, codebook_size=K, commitment_loss_weight=beta, codebook_loss_weight=w_cb,
                     l2_normalized=bool(l2), stochastic=bool(stoch), stochastic_temperature=T,
                     same_index_shape=bool(same))
    return cfg, bool(eval_det)


def rel_err(a, b):
    a = torch.as_tensor(a).double().flatten()
    b = torch.as_tensor(b).double().flatten()
    den = b.norm().item()
    if den == 0:
        return (a - b).norm().item()
    return ((a - b).norm() / den).item()


def assert_close(a, b, what, tol=REL_TOL):
    e = rel_err(a, b)
    assert e <= tol, f"{what}: relative error {e:.3e} > {tol:.1e}"


def assert_indices_match(idx, ref_idx, ref_scores, larger_is_better, what="indices"):
    """Bit-exact, except rows where the code we picked scores within GAP_TOL of the
    reference's best in the reference's own ranking matrix."""
    idx = torch.as_tensor(idx).reshape(-1).long().cpu()
    ref_idx = torch.as_tensor(ref_idx).reshape(-1).long().cpu()
    bad = (idx != ref_idx).nonzero().flatten()
    if bad.numel() == 0:
        return 0
    gaps = O.gap_to_choice(ref_scores[bad], idx[bad], larger_is_better)
    worst = float(gaps.max())
    assert worst < GAP_TOL, (f"{what}: {bad.numel()} rows differ and the worst has a reference score gap "
                             f"{worst:.3e} >= {GAP_TOL:.0e} (row {int(bad[gaps.argmax()])})")
    return int(bad.numel())

"""Generate tests/golden/*.npz by running the UNMODIFIED reference module on CPU.

TEST INFRASTRUCTURE.  Run in the build container (needs /root/reference):

    python oracle/make_golden.py

Every fixture stores the seeded inputs and what the reference's own
``Bottleneck`` / ``SimpleVectorQuantizer`` (models/bottleneck.py) returned for them:
indices, tensors, losses, autograd gradients and the per-row top-2 gap of the
reference's ranking matrix.  torch version and thread count are recorded.
"""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle.ref_loader import load_reference, reference_root  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def _np(t):
    t = t.detach()
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy().copy()  # raw bits
    return t.numpy().copy()


def _ranking(ref_vq, z):
    """The matrix the reference ranks by, recomputed with its own expressions."""
    import torch.nn.functional as F
    with torch.no_grad():
        zf = z.float()
        if ref_vq.l2_normalized:
            zf = F.normalize(zf, p=2, dim=-1)
        emb = ref_vq.get_emb()
        zf = zf.reshape(-1, zf.shape[-1])
        if ref_vq.stochastic:
            return torch.einsum("bd,nd->bn", zf, emb), True
        d = (torch.sum(zf ** 2, dim=1, keepdim=True) + torch.sum(emb ** 2, dim=1)
             - 2 * torch.einsum("bd,dn->bn", zf, emb.t()))
        return d, False


def run_vq(bott, name, *, B, n, D, K, seed, l2=True, stochastic=False, eval_det=False, T=0.03,
           z_dtype=torch.float32, z_override=None, w_override=None, same_index_shape=True,
           beta=0.25, w_cb=1.0, store_tensors=True, g_lossq=0.7):
    torch.manual_seed(seed)
    vq = bott.SimpleVectorQuantizer(dim=D, codebook_size=K, commitment_loss_weight=beta,
                                    codebook_loss_weight=w_cb, l2_normalized=l2,
                                    stochastic=stochastic, stochastic_temperature=T,
                                    same_index_shape=same_index_shape)
    if w_override is not None:
        with torch.no_grad():
            vq.embedding.weight.copy_(w_override)
    z = torch.randn(B, n, D) if z_override is None else z_override.clone()
    z = z.to(z_dtype)
    G = torch.randn(B, n, D)
    if eval_det:
        vq.eval()
        vq.set_eval_deterministic(True)
    zr = z.clone().requires_grad_(True)
    out = vq(zr)
    obj = (out["regularized_z"] * G).sum() + g_lossq * out["loss_q"]
    obj.backward()
    scores, larger = _ranking(vq, z)
    k2 = min(2, scores.shape[1])
    top = torch.topk(scores, k2, dim=1, largest=larger).values
    gap = (top[:, 0] - top[:, -1]).abs() if k2 == 2 else torch.full((scores.shape[0],), np.inf)
    rec = dict(
        z=_np(z), z_is_bf16=np.array(z_dtype == torch.bfloat16), weight=_np(vq.embedding.weight),
        G=_np(G), g_lossq=np.float32(g_lossq),
        idx=_np(out["bottleneck_rep"]).astype(np.int64), top2_gap=_np(gap),
        loss_q=_np(out["loss_q"]), loss_commit=_np(out["loss_commit"]), loss_codebook=_np(out["loss_codebook"]),
        cfg=np.array([D, K, int(l2), int(stochastic), int(eval_det), int(same_index_shape)], dtype=np.int64),
        cfg_f=np.array([beta, w_cb, T], dtype=np.float64),
    )
    if store_tensors:
        rec.update(unregularized_z=_np(out["unregularized_z"]), regularized_z=_np(out["regularized_z"]),
                   emb=_np(out["emb"]), grad_z=_np(zr.grad.float()), grad_weight=_np(vq.embedding.weight.grad))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(f"{name}: N={B*n} K={K} D={D} min_gap={float(gap.min()):.3e} loss_q={float(out['loss_q']):.6f}")


def run_bottleneck(bott, registry, name, *, B, n, C, D, K, seed, stochastic=False, eval_det=False,
                   strided=False):
    torch.manual_seed(seed)
    spec = {"name": "bottleneck", "args": {
        "bottleneck_dim": D, "norm": "none",
        "regularizer": {"name": "vq", "args": {
            "codebook_size": K, "commitment_loss_weight": 0.25, "codebook_loss_weight": 1.0,
            "entropy_loss_weight": 0.0, "entropy_loss_temperature": 0.01, "l2_normalized": True,
            "stochastic": stochastic, "stochastic_temperature": 0.03}}}}
    m = registry.make(spec, args={"token_nums": n, "input_dim": C, "output_dim": C})
    if eval_det:
        m.eval()
        m.regularizer.set_eval_deterministic(True)
    if strided:  # the encoder hands over h[:, -n:, :] (models/transformer.py:69): batch-strided
        big = torch.randn(B, 2 * n, C)
        x = big[:, -n:, :]
    else:
        x = torch.randn(B, n, C)
    Gout = torch.randn(B, n, C)
    xr = x.detach().clone().requires_grad_(True) if not strided else x.detach().requires_grad_(True)
    out = m(xr)
    obj = (out["output"] * Gout).sum() + 0.1 * out["loss_q"]
    obj.backward()
    keys = sorted(out.keys())
    rec = dict(
        x=_np(x.contiguous()), Gout=_np(Gout), w_in=_np(m.in_linear.weight), b_in=_np(m.in_linear.bias),
        w_out=_np(m.out_linear.weight), b_out=_np(m.out_linear.bias), weight=_np(m.regularizer.embedding.weight),
        output=_np(out["output"]), projected_z=_np(out["projected_z"]), idx=_np(out["bottleneck_rep"]),
        regularized_z=_np(out["regularized_z"]), unregularized_z=_np(out["unregularized_z"]),
        loss_q=_np(out["loss_q"]), loss_commit=_np(out["loss_commit"]),
        input_norm_first=np.float64(out["input_norm_first"]), input_norm_last=np.float64(out["input_norm_last"]),
        grad_x=_np(xr.grad), grad_w_in=_np(m.in_linear.weight.grad), grad_b_in=_np(m.in_linear.bias.grad),
        grad_w_out=_np(m.out_linear.weight.grad), grad_b_out=_np(m.out_linear.bias.grad),
        grad_weight=_np(m.regularizer.embedding.weight.grad),
        keys=np.array(keys), state_keys=np.array(sorted(m.state_dict().keys())),
        cfg=np.array([B, n, C, D, K, int(stochastic), int(eval_det)], dtype=np.int64),
    )
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(f"{name}: keys={keys}")


def run_config1(bott, registry):
    """BASELINE configs[0]: bottleneck forward on CPU, deterministic argmax, B=4x1024,
    codebook per cfgs/larp_tokenizer.yaml:39-54 (D=24, K=8192, C=768).  Stored at the
    quantizer boundary (z, codebook, indices, losses, gaps) to keep the fixture small."""
    for tag, stochastic, eval_det in (("argmin", False, False), ("evaldet", True, True)):
        torch.manual_seed(0)
        spec = {"name": "bottleneck", "args": {
            "bottleneck_dim": 24, "norm": "none",
            "regularizer": {"name": "vq", "args": {
                "codebook_size": 8192, "commitment_loss_weight": 0.25, "codebook_loss_weight": 1.0,
                "entropy_loss_weight": 0.0, "entropy_loss_temperature": 0.01, "l2_normalized": True,
                "stochastic": stochastic, "stochastic_temperature": 0.03}}}}
        m = registry.make(spec, args={"token_nums": 1024, "input_dim": 768, "output_dim": 768})
        if eval_det:
            m.eval()
            m.regularizer.set_eval_deterministic(True)
        x = torch.randn(4, 1024, 768)
        with torch.no_grad():
            out = m(x)
        z = out["projected_z"]
        scores, larger = _ranking(m.regularizer, z)
        top = torch.topk(scores, 2, dim=1, largest=larger).values
        gap = (top[:, 0] - top[:, 1]).abs()
        # fp16 storage of x would change it; store only a checksum of x and the exact z
        rec = dict(z=_np(z), weight=_np(m.regularizer.embedding.weight),
                   idx=_np(out["bottleneck_rep"]).astype(np.int16),
                   top2_gap=_np(gap), loss_q=_np(out["loss_q"]), loss_commit=_np(out["loss_commit"]),
                   x_sum=np.float64(x.double().sum()), output_sum=np.float64(out["output"].double().sum()),
                   output_abs_sum=np.float64(out["output"].double().abs().sum()),
                   input_norm_first=np.float64(out["input_norm_first"]),
                   cfg=np.array([24, 8192, 1, int(stochastic), int(eval_det), 1], dtype=np.int64),
                   cfg_f=np.array([0.25, 1.0, 0.03], dtype=np.float64))
        np.savez_compressed(os.path.join(OUT, f"config1_{tag}.npz"), **rec)
        print(f"config1_{tag}: min_gap={float(gap.min()):.3e} used_codes={len(torch.unique(out['bottleneck_rep']))}")


def run_stoch_probs(bott):
    """Reference softmax probabilities for the chi-square test (bottleneck.py:275-276)."""
    import torch.nn.functional as F
    torch.manual_seed(11)
    for D, K in ((16, 512), (24, 8192)):
        vq = bott.SimpleVectorQuantizer(dim=D, codebook_size=K, l2_normalized=True, stochastic=True,
                                        stochastic_temperature=0.03)
        z = torch.randn(1, 8, D)
        with torch.no_grad():
            zf = F.normalize(z.float(), p=2, dim=-1).reshape(-1, D)
            cos = torch.einsum("bd,nd->bn", zf, vq.get_emb())
            probs = F.softmax(cos * vq.stochastic_temperature_inv, dim=-1)
            # expected commitment loss under the reference distribution (unit vectors: |q-z|^2 = 2-2cos)
            exp_loss = (probs.double() * (2 - 2 * cos.double())).sum(1).mean() / D
        np.savez_compressed(os.path.join(OUT, f"stoch_probs_d{D}_k{K}.npz"), z=_np(z),
                            weight=_np(vq.embedding.weight), probs=_np(probs), cos=_np(cos),
                            expected_loss_commit=np.float64(exp_loss), t_inv=np.float64(vq.stochastic_temperature_inv))
        print(f"stoch_probs_d{D}_k{K}: max-prob mean={float(probs.max(1).values.mean()):.3f}")


def run_stats():
    """Trainer code-usage statistics (trainers/larp_tokenizer_trainer.py:335-352) with the
    reference's utils/statistics.py loaded from its file."""
    import torch.nn.functional as F
    spec = importlib.util.spec_from_file_location("ref_statistics",
                                                  os.path.join(reference_root(), "utils", "statistics.py"))
    st = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(st)
    torch.manual_seed(5)
    K = 1024
    rep = torch.randint(0, 300, (4, 256))
    rep[1:] = torch.randint(0, K, (3, 256))
    used = rep[0].flatten()
    counts = torch.bincount(used, minlength=K)
    counts_b = torch.bincount(rep.flatten(), minlength=K)
    enc = F.one_hot(used, K).float().reshape(-1, K)
    avg = enc.mean(0)
    rec = dict(rep=_np(rep), K=np.int64(K),
               index_usage=np.float64(st.index_usage_percentage(counts)),
               index_usage_batch=np.float64(st.index_usage_percentage(counts_b)),
               kl_uni=np.float64(st.kl_divergence_from_uniform(counts)),
               perplexity=np.float64((-(avg * torch.log(avg + 1e-10)).sum()).exp().item()))
    np.savez_compressed(os.path.join(OUT, "code_usage_stats.npz"), **rec)
    print("code_usage_stats:", {k: float(v) for k, v in rec.items() if k not in ("rep", "K")})


def _load_ref_fsq():
    """the UNMODIFIED reference models/model_new/quantizer/fsq.py (needs only torch, einops, numpy)"""
    import importlib.util
    import sys
    from oracle.ref_loader import reference_root
    path = os.path.join(reference_root(), "models", "model_new", "quantizer", "fsq.py")
    spec = importlib.util.spec_from_file_location("ref_fsq", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_fsq"] = mod
    spec.loader.exec_module(mod)
    return mod


def run_sq(fsq, name, *, shape, D, K, seed, l2, beta=0.25, input_format="blc", frozen=False):
    """`VectorQuantizer` ('sq' bottleneck): outputs + grads from the reference class itself"""
    import tempfile
    torch.manual_seed(seed)
    W = torch.randn(K, D)
    if l2:
        W = torch.nn.functional.normalize(W, dim=-1)      # like the pre-normalised Leech-lattice table
    W[K // 3] = W[K // 7]                                  # exact tie -> lowest index
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "cb.npy")
        np.save(path, W.numpy())
        q = fsq.VectorQuantizer(n_embed=K, embed_dim=D, l2_norm=l2, beta=beta, input_format=input_format,
                                predefined_codebook=path, freeze_codebook=frozen)
    q.train()
    z = torch.randn(*shape, requires_grad=True)
    out = q(z)
    G = torch.randn_like(out["output"])
    ((out["output"] * G).sum() + 0.7 * out["loss_codebook"]).backward()
    rec = {"z": _np(z), "weight": _np(W), "G": _np(G), "output": _np(out["output"]), "loss_codebook": _np(out["loss_codebook"]),
           "grad_z": _np(z.grad), "l2": np.array(l2), "beta": np.array(beta), "frozen": np.array(frozen),
           "input_format": np.array(input_format)}
    if not frozen:
        rec["grad_weight"] = _np(q.embedding.weight.grad)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print("wrote", name)


def run_entropy(bott, name, *, B, n, D, K, seed, l2, w_ent, T_ent, z_scale=1.0):
    """`SimpleVectorQuantizer` with entropy_loss_weight > 0 (bottleneck.py:300-304 -> entropy_loss :12-33):
    outputs, the three entropy scalars and grads from the reference class itself.  The objective puts separate
    upstream gradients on loss_q, per_sample_entropy and codebook_entropy."""
    torch.manual_seed(seed)
    vq = bott.SimpleVectorQuantizer(dim=D, codebook_size=K, l2_normalized=l2, entropy_loss_weight=w_ent,
                                    entropy_loss_temperature=T_ent)
    z = torch.randn(B, n, D) * z_scale
    G = torch.randn(B, n, D)
    zr = z.clone().requires_grad_(True)
    out = vq(zr)
    obj = ((out["regularized_z"] * G).sum() + 0.7 * out["loss_q"] + 0.3 * out["per_sample_entropy"]
           - 0.2 * out["codebook_entropy"])
    obj.backward()
    rec = dict(z=_np(z), weight=_np(vq.embedding.weight), G=_np(G), idx=_np(out["bottleneck_rep"]).astype(np.int64),
               cfg=np.array([D, K, int(l2)], dtype=np.int64), cfg_f=np.array([w_ent, T_ent], dtype=np.float64),
               grad_z=_np(zr.grad), grad_weight=_np(vq.embedding.weight.grad))
    for k in ("loss_q", "loss_commit", "loss_codebook", "loss_entropy", "per_sample_entropy", "codebook_entropy"):
        rec[k] = _np(out[k])
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(f"{name}: loss_entropy={float(out['loss_entropy']):.6f} sample={float(out['per_sample_entropy']):.6f} "
          f"avg={float(out['codebook_entropy']):.6f}")


def _load_ref_cosmos():
    """the UNMODIFIED reference models/cosmos.py under a stub `models` package that only provides `register`"""
    import importlib.util
    import sys
    import types
    from oracle.ref_loader import reference_root
    saved = sys.modules.get("models")
    shell = types.ModuleType("models")
    shell.__path__ = []
    shell.register = lambda name: (lambda cls: cls)
    sys.modules["models"] = shell
    try:
        spec = importlib.util.spec_from_file_location("ref_cosmos", os.path.join(reference_root(), "models", "cosmos.py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules["ref_cosmos"] = mod
        spec.loader.exec_module(mod)
    finally:
        if saved is None:
            sys.modules.pop("models", None)
        else:
            sys.modules["models"] = saved
    return mod


def run_simvq(cosmos, name, *, shape, K, seed, beta=0.25, legacy=True):
    """`SimVQ` (models/cosmos.py:772-876): outputs + grads from the reference class itself"""
    torch.manual_seed(seed)
    D = shape[1]
    q = cosmos.SimVQ(n_e=K, e_dim=D, beta=beta, legacy=legacy)
    with torch.no_grad():
        q.embedding.weight[K // 3] = q.embedding.weight[K // 7]     # exact tie -> lowest index
    z = (torch.randn(*shape) * 0.5).requires_grad_(True)
    z_q, loss, (_, ind) = q(z)
    G = torch.randn_like(z_q)
    ((z_q * G).sum() + 0.7 * loss).backward()
    assert q.embedding.weight.grad is None
    rec = {"z": _np(z), "anchors": _np(q.embedding.weight), "proj_w": _np(q.embedding_proj.weight),
           "proj_b": _np(q.embedding_proj.bias), "G": _np(G), "z_q": _np(z_q), "loss": _np(loss),
           "indices": _np(ind).astype(np.int64), "grad_z": _np(z.grad), "grad_proj_w": _np(q.embedding_proj.weight.grad),
           "grad_proj_b": _np(q.embedding_proj.bias.grad), "beta": np.array(beta), "legacy": np.array(legacy)}
    idx2 = ind[:1]
    rec["entry_indices"] = _np(idx2).astype(np.int64)
    rec["entry"] = _np(q.get_codebook_entry(idx2))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print("wrote", name, "loss", float(loss))


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    bott, registry = load_reference()
    if os.environ.get("GOLDEN_ONLY") == "entropy":
        run_entropy(bott, "ent_l2_d16_k300", B=2, n=70, D=16, K=300, seed=31, l2=True, w_ent=0.1, T_ent=0.01)
        run_entropy(bott, "ent_l2_d24_k1000_warm", B=3, n=100, D=24, K=1000, seed=32, l2=True, w_ent=0.05, T_ent=0.1)
        run_entropy(bott, "ent_nol2_d8_k130", B=1, n=129, D=8, K=130, seed=33, l2=False, w_ent=0.2, T_ent=1.0, z_scale=0.3)
        run_entropy(bott, "ent_l2_d32_k64", B=2, n=33, D=32, K=64, seed=34, l2=True, w_ent=0.1, T_ent=0.02)
        return
    if os.environ.get("GOLDEN_ONLY") == "simvq":
        cosmos = _load_ref_cosmos()
        run_simvq(cosmos, "simvq_image_d16_k300", shape=(2, 16, 5, 6), K=300, seed=41)
        run_simvq(cosmos, "simvq_video_d24_k1000_nonlegacy", shape=(2, 24, 3, 4, 5), K=1000, seed=42, legacy=False)
        run_simvq(cosmos, "simvq_video_d320_k200", shape=(1, 320, 2, 3, 3), K=200, seed=43)
        run_simvq(cosmos, "simvq_image_d40_k129", shape=(3, 40, 4, 4), K=129, seed=44, beta=0.5)
        return
    if os.environ.get("GOLDEN_ONLY") == "sq":
        fsq = _load_ref_fsq()
        run_sq(fsq, "sq_l2_d24_k1000_blc", shape=(2, 48, 24), D=24, K=1000, seed=21, l2=True)
        run_sq(fsq, "sq_l2_d24_k4096_frozen", shape=(3, 64, 24), D=24, K=4096, seed=22, l2=True, frozen=True)
        run_sq(fsq, "sq_nol2_d16_k300_bchw", shape=(2, 16, 5, 6), D=16, K=300, seed=23, l2=False, input_format="bchw")
        return
    run_vq(bott, "vq_argmin_d16_k500", B=2, n=96, D=16, K=500, seed=1)
    run_vq(bott, "vq_argmin_d24_k1024", B=3, n=64, D=24, K=1024, seed=2)
    run_vq(bott, "vq_argmin_d8_k64", B=1, n=33, D=8, K=64, seed=3)
    run_vq(bott, "vq_argmin_d32_k777", B=2, n=50, D=32, K=777, seed=4)
    run_vq(bott, "vq_argmin_d64_k256", B=2, n=40, D=64, K=256, seed=5)
    run_vq(bott, "vq_argmin_n1", B=1, n=1, D=16, K=128, seed=6)
    run_vq(bott, "vq_evaldet_d16_k512", B=2, n=64, D=16, K=512, seed=7, stochastic=True, eval_det=True)
    run_vq(bott, "vq_nol2_d16_k300", B=2, n=64, D=16, K=300, seed=8, l2=False)
    run_vq(bott, "vq_bf16_d16_k512", B=2, n=64, D=16, K=512, seed=9, z_dtype=torch.bfloat16)
    run_vq(bott, "vq_flatidx_d16_k128", B=2, n=16, D=16, K=128, seed=10, same_index_shape=False)
    # edge cases: duplicate code rows (ties -> lowest index), zero z row, zero code row
    torch.manual_seed(12)
    W = torch.randn(64, 16) * 0.3
    W[40] = W[7]
    W[41] = W[7]
    W[13] = 0.0
    z = torch.randn(1, 24, 16)
    z[0, 3] = 0.0
    z[0, 5] = W[7] * 2.5  # lands exactly on the duplicated direction
    z[0, 6] = W[41]
    run_vq(bott, "vq_edge_ties_zero", B=1, n=24, D=16, K=64, seed=12, z_override=z, w_override=W)
    run_vq(bott, "vq_edge_ties_zero_evaldet", B=1, n=24, D=16, K=64, seed=12, z_override=z, w_override=W,
           stochastic=True, eval_det=True)
    run_bottleneck(bott, registry, "bottleneck_small", B=2, n=32, C=64, D=16, K=256, seed=21)
    run_bottleneck(bott, registry, "bottleneck_strided_evaldet", B=2, n=24, C=48, D=24, K=200, seed=22,
                   stochastic=True, eval_det=True, strided=True)
    run_config1(bott, registry)
    run_stoch_probs(bott)
    run_stats()
    with open(os.path.join(OUT, "PROVENANCE.txt"), "w") as f:
        f.write(f"generated by oracle/make_golden.py from the unmodified reference at {reference_root()}\n"
                f"torch {torch.__version__}, CPU fp32, threads={torch.get_num_threads()}\n"
                "reference files executed: models/models.py, models/bottleneck.py, utils/statistics.py\n")


if __name__ == "__main__":
    main()

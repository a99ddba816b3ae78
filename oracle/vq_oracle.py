"""CPU oracle for the LARP quantization bottleneck.  TEST INFRASTRUCTURE, NOT PRODUCT.

A restatement, in plain fp32 torch-CPU tensor ops, of the algorithm in the reference's
``models/bottleneck.py`` (``SimpleVectorQuantizer`` :203-344, ``Bottleneck`` :65-188,
``entropy_loss`` :12-33) and of the trainer's code-usage statistics
(``trainers/larp_tokenizer_trainer.py:335-352``, ``utils/statistics.py:5-32``).

The reference is itself torch code, so the restatement uses the same primitive ops
(normalize, einsum, argmin, softmax, multinomial, embedding) in the same order: on the
same CPU and torch build its results are bit-identical to the reference, which is what
the "index mismatch only where the reference top-2 gap < 1e-6" rule needs.

Parity pin: ``tests/golden/*.npz`` were produced by importing the UNMODIFIED reference
module in the build container (``oracle/make_golden.py`` + ``oracle/ref_loader.py``);
``tests/test_oracle_golden.py`` checks this file against every one of them.  The
reference has no tests/golden vectors of its own.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn.functional as F

EPS = 1e-12  # F.normalize default eps (reference bottleneck.py:255,267)


@dataclass
class VQConfig:
    """Constructor arguments of the reference quantizer (bottleneck.py:205-217)."""
    dim: int
    codebook_size: int
    commitment_loss_weight: float = 0.25
    entropy_loss_weight: float = 0.0
    entropy_loss_temperature: float = 0.01
    l2_normalized: bool = False
    same_index_shape: bool = True
    stochastic: bool = False
    stochastic_temperature: float = 1.0
    codebook_loss_weight: float = 1.0


def codebook_view(weight: torch.Tensor, l2_normalized: bool) -> torch.Tensor:
    """get_emb (bottleneck.py:252-259): the codebook as the quantizer sees it."""
    assert weight.dtype == torch.float32
    return F.normalize(weight, p=2, dim=-1) if l2_normalized else weight


def entropy_terms(affinity: torch.Tensor, loss_type: str = "softmax", temperature: float = 0.01):
    """entropy_loss (bottleneck.py:12-33).  NOTE: divides its argument in place, as the
    reference does, and adds 1e-5 to the logits before log_softmax."""
    flat = affinity.view(-1, affinity.shape[-1])
    flat /= temperature
    probs = F.softmax(flat, dim=-1)
    log_probs = F.log_softmax(flat + 1e-5, dim=-1)
    if loss_type == "softmax":
        target = probs
    elif loss_type == "argmax":
        codes = torch.argmax(flat, dim=-1)
        onehots = F.one_hot(codes, num_classes=flat.shape[-1]).to(flat.dtype)
        target = probs - (probs - onehots).detach()
    else:
        raise ValueError(f"Entropy loss {loss_type} not supported")
    avg_probs = torch.mean(target, dim=0)
    avg_entropy = -torch.sum(avg_probs * torch.log(avg_probs + 1e-5))
    sample_entropy = -torch.mean(torch.sum(target * log_probs, dim=-1))
    return sample_entropy - avg_entropy, sample_entropy, avg_entropy


def entropy_closed_form(z_flat: torch.Tensor, emb: torch.Tensor, temperature: float, g_loss: float = 1.0,
                        g_sample: float = 0.0, g_avg: float = 0.0):
    """The streaming formulation the CUDA passes implement (csrc/vq_entropy.cuh), in float64 with the full matrix:
    the three scalars of `entropy_loss(-d, temperature)` (bottleneck.py:12-33, "softmax" type) and the closed-form
    gradients wrt the latent rows and the code rows for upstream gradients (g_loss, g_sample, g_avg) on
    (loss, sample_entropy, avg_entropy).  Referees autograd of `entropy_terms` in tests/test_entropy.py."""
    z = z_flat.double()
    e = emb.double()
    N = z.shape[0]
    a = (2 * z @ e.t() - (z ** 2).sum(1, keepdim=True) - (e ** 2).sum(1)) / temperature
    lse = torch.logsumexp(a, dim=1, keepdim=True)
    logp = a - lse
    p = logp.exp()
    ent = (p * logp).sum(1, keepdim=True)                    # sum_k p log p per row (= -H_i)
    pbar = p.mean(0)
    sample = -ent.mean()
    avg = -(pbar * torch.log(pbar + 1e-5)).sum()
    w = -(torch.log(pbar + 1e-5) + pbar / (pbar + 1e-5))     # d avg / d pbar_k
    c = (p * w).sum(1, keepdim=True)
    cS, cA = g_loss + g_sample, g_avg - g_loss
    g = p / N * (-cS * (logp - ent) + cA * (w - c))          # d objective / d a_ik
    gz = (2 / temperature) * (g @ e - z * g.sum(1, keepdim=True))
    ge = (2 / temperature) * (g.t() @ z - e * g.sum(0).unsqueeze(1))
    return (sample - avg, sample, avg), gz, ge


def vq_scores(z_flat: torch.Tensor, emb: torch.Tensor, stochastic: bool) -> torch.Tensor:
    """The N x K matrix the reference ranks codes by.

    stochastic=False: squared distance d (bottleneck.py:282-289), smaller is better.
    stochastic=True : cosine similarity (bottleneck.py:275), larger is better.
    """
    if stochastic:
        return torch.einsum("bd,nd->bn", z_flat, emb)
    return (
        torch.sum(z_flat ** 2, dim=1, keepdim=True)
        + torch.sum(emb ** 2, dim=1)
        - 2 * torch.einsum("bd,dn->bn", z_flat, emb.t())
    )


def vq_forward(
    z: torch.Tensor,
    weight: torch.Tensor,
    cfg: VQConfig,
    *,
    training: bool = True,
    eval_deterministic: bool = False,
    temperature_inv: Optional[float] = None,
    generator: Optional[torch.Generator] = None,
    forced_indices: Optional[torch.Tensor] = None,
    return_scores: bool = False,
) -> dict:
    """SimpleVectorQuantizer.forward (bottleneck.py:262-324) on CPU fp32.

    ``z`` (B, n, D) any float dtype; ``weight`` (K, D) fp32 (may require grad).
    ``forced_indices`` (N,) replaces the selection step: used to push the kernel's own
    stochastic draws through the reference's gather/loss/STE arithmetic.
    Returns the reference's 10-key dict (+ 'scores', 'top2_gap' when asked).
    """
    z = z.float()
    assert z.dim() == 3, "Input shape must be (batch, n_tokens, e_dim)"
    if cfg.l2_normalized:
        z = F.normalize(z, p=2, dim=-1)
    emb = codebook_view(weight, cfg.l2_normalized)
    z_flat = z.reshape(-1, z.shape[-1])
    t_inv = temperature_inv if temperature_inv is not None else (
        1.0 / cfg.stochastic_temperature if cfg.stochastic_temperature > 0 else 10.0)

    scores = None
    d = None
    if cfg.stochastic:
        assert cfg.l2_normalized, "Stochastic sampling requires l2 normalization"
        scores = vq_scores(z_flat, emb, True)
        probs = F.softmax(scores * t_inv, dim=-1)
        if forced_indices is not None:
            idx = forced_indices.reshape(-1).long()
        elif eval_deterministic and not training:
            idx = torch.argmax(probs, dim=-1)
        else:
            idx = torch.multinomial(probs, 1, generator=generator).squeeze(-1)
    else:
        d = vq_scores(z_flat, emb, False)
        scores = d
        idx = forced_indices.reshape(-1).long() if forced_indices is not None else torch.argmin(d, dim=1)

    quantized = F.embedding(idx, emb).view(z.shape)
    loss_commit = ((quantized.detach() - z) ** 2).mean()
    loss_codebook = ((quantized - z.detach()) ** 2).mean()
    if cfg.entropy_loss_weight > 0:
        if d is None:
            # the reference hits UnboundLocalError here (bottleneck.py:301 with :272-281)
            raise UnboundLocalError("cannot access local variable 'd' where it is not associated with a value")
        loss_entropy, sample_entropy, avg_entropy = entropy_terms(-d, temperature=cfg.entropy_loss_temperature)
    else:
        loss_entropy = sample_entropy = avg_entropy = torch.tensor(0.0, dtype=z.dtype)
    loss = (cfg.commitment_loss_weight * loss_commit + cfg.codebook_loss_weight * loss_codebook
            + cfg.entropy_loss_weight * loss_entropy)
    ste = z + (quantized - z).detach()
    rep = idx.reshape(ste.shape[0], ste.shape[1]) if cfg.same_index_shape else idx
    out = {
        "unregularized_z": z,
        "emb": emb,
        "regularized_z": ste,
        "bottleneck_rep": rep,
        "loss_q": loss,
        "loss_commit": loss_commit,
        "loss_codebook": loss_codebook,
        "loss_entropy": loss_entropy,
        "per_sample_entropy": sample_entropy,
        "codebook_entropy": avg_entropy,
    }
    if return_scores:
        with torch.no_grad():
            sc = scores.detach()
            out["scores"] = sc
            out["top2_gap"] = top2_gap(sc, larger_is_better=cfg.stochastic)
    return out


def top2_gap(scores: torch.Tensor, larger_is_better: bool) -> torch.Tensor:
    """Per-row |best - second best| of the reference's own ranking matrix."""
    if scores.shape[1] < 2:
        return torch.full((scores.shape[0],), float("inf"))
    top = torch.topk(scores, 2, dim=1, largest=larger_is_better).values
    return (top[:, 0] - top[:, 1]).abs()


def gap_to_choice(scores: torch.Tensor, idx: torch.Tensor, larger_is_better: bool) -> torch.Tensor:
    """|reference best score - reference score of the code ``idx`` picked| per row.
    A kernel index differing from the reference's is acceptable only if this is < 1e-6
    (BASELINE north_star)."""
    best = scores.max(dim=1).values if larger_is_better else scores.min(dim=1).values
    mine = scores.gather(1, idx.reshape(-1, 1).long()).squeeze(1)
    return (best - mine).abs()


def codebook_lookup(indices: torch.Tensor, weight: torch.Tensor, l2_normalized: bool, shape=None) -> torch.Tensor:
    """get_codebook_entry / decode (bottleneck.py:327-344)."""
    flat = indices.reshape(-1)
    z_q = F.embedding(flat, weight)
    if l2_normalized:
        z_q = F.normalize(z_q, p=2, dim=-1)
    return z_q.reshape(shape) if shape is not None else z_q.reshape([*indices.shape, weight.shape[1]])


def apply_norm(z, norm: Optional[str], norm_layer):
    """Bottleneck.project_in's norm dispatch (bottleneck.py:143-160).  'ln_d_na' falls
    through every branch in the reference (constructed at :125-126, never applied)."""
    if norm is None:
        return z
    z = z.float()
    if norm == "bn_bn":
        return norm_layer(z.transpose(1, 2)).transpose(1, 2)
    if norm == "bn_b":
        b, n, d = z.shape
        return norm_layer(z.reshape(b, n * d)).reshape(b, n, d)
    if norm in ("ln_nd", "ln_d"):
        return norm_layer(z)
    return z


def bottleneck_forward(
    x: torch.Tensor,
    w_in: torch.Tensor, b_in: torch.Tensor,
    w_out: torch.Tensor, b_out: torch.Tensor,
    weight: torch.Tensor,
    cfg: VQConfig,
    *, norm: Optional[str] = None, norm_layer=None, **vq_kwargs,
) -> dict:
    """Bottleneck.forward (bottleneck.py:170-188): diagnostics, in_linear, VQ, out_linear."""
    assert x.dim() == 3, "Input shape must be (batch, n_tokens, e_dim)"
    input_norm_first = torch.norm(x[:, 0, :], dim=-1).mean().item()
    input_norm_last = torch.norm(x[:, -1, :], dim=-1).mean().item()
    z = F.linear(x, w_in, b_in)
    z = apply_norm(z, norm, norm_layer)
    reg = vq_forward(z, weight, cfg, **vq_kwargs)
    x_hat = F.linear(reg["regularized_z"], w_out, b_out)
    rep = reg.pop("bottleneck_rep")
    return {"output": x_hat, "bottleneck_rep": rep, "projected_z": z,
            "input_norm_first": input_norm_first, "input_norm_last": input_norm_last, **reg}


def bottleneck_decode(indices, weight, l2_normalized, w_out, b_out):
    """Bottleneck.decode (bottleneck.py:166-168)."""
    return F.linear(codebook_lookup(indices, weight, l2_normalized), w_out, b_out)


def sq_forward(z: torch.Tensor, weight: torch.Tensor, l2_norm: bool, beta: float, input_format: str = "blc") -> dict:
    """`VectorQuantizer.forward` of the sibling quantizer family (models/model_new/quantizer/fsq.py:176-214;
    built as bottleneck_type 'sq' with K=196 560, D=24, models/larp_tokenizer.py:230-233): nearest code by
    -cos (l2) or squared distance, loss = beta*mean(sum_d (q.detach()-z)^2) + mean(sum_d (q-z.detach())^2)
    (:207: summed over the code dimension, then averaged), straight-through output (:209).  Also returns
    the indices and the usage entropy the reference computes and drops (:195-196, :216-219)."""
    if input_format == "bchw":
        z = z.permute(0, 2, 3, 1)
    D = weight.shape[1]
    if l2_norm:
        z = F.normalize(z, dim=-1)
        z_flat = z.reshape(-1, D)
        d = -z_flat @ F.normalize(weight, dim=-1).t()
    else:
        z_flat = z.reshape(-1, D)
        d = torch.sum(z_flat ** 2, dim=1, keepdim=True) + torch.sum(weight ** 2, dim=1) - 2 * z_flat @ weight.t()
    idx = torch.argmin(d.detach(), dim=1)
    count = torch.bincount(idx, minlength=weight.shape[0]).float()
    probs = (count + 1e-4) / (count + 1e-4).sum()
    entropy = -(probs * torch.log(probs)).sum()
    z_q = F.embedding(idx, weight).view(z.shape)
    if l2_norm:
        z_q = F.normalize(z_q, dim=-1)
    loss = beta * torch.mean(((z_q.detach() - z) ** 2).sum(dim=-1)) + torch.mean(((z_q - z.detach()) ** 2).sum(dim=-1))
    out = z + (z_q - z).detach()
    if input_format == "bchw":
        out = out.permute(0, 3, 1, 2)
    return {"output": out, "loss_codebook": loss, "indices": idx, "cb_entropy": entropy, "scores": d}


def simvq_forward(z: torch.Tensor, anchors: torch.Tensor, proj_w: torch.Tensor, proj_b: torch.Tensor, beta: float,
                  legacy: bool = True) -> dict:
    """`SimVQ.forward` (models/cosmos.py:792-855): frozen anchor table through one Linear (:810), nearest projected
    code by squared distance (:814-819), legacy / non-legacy loss weighting (:826-831), straight-through output
    (:835), channel-first in and out (:795-803, :838-847).  z: (B,C,H,W) or (B,C,T,H,W)."""
    video = z.dim() == 5
    z_cl = (z.permute(0, 2, 3, 4, 1) if video else z.permute(0, 2, 3, 1)).contiguous()
    D = anchors.shape[1]
    z_flat = z_cl.view(-1, D)
    table = F.linear(anchors, proj_w, proj_b)
    d = torch.sum(z_flat ** 2, dim=1, keepdim=True) + torch.sum(table ** 2, dim=1) - 2 * z_flat @ table.t()
    idx = torch.argmin(d, dim=1)
    z_q = F.embedding(idx, table).view(z_cl.shape)
    commit = torch.mean((z_q.detach() - z_cl) ** 2)
    codebook = torch.mean((z_q - z_cl.detach()) ** 2)
    loss = commit + beta * codebook if legacy else beta * commit + codebook
    z_q = z_cl + (z_q - z_cl).detach()
    if video:
        z_q = z_q.permute(0, 4, 1, 2, 3).contiguous()
        ind = idx.view(z.shape[0], z.shape[2], z.shape[3], z.shape[4])
    else:
        z_q = z_q.permute(0, 3, 1, 2).contiguous()
        ind = idx.view(z.shape[0], z.shape[2], z.shape[3])
    return {"z_q": z_q, "loss": loss, "indices": ind, "scores": d.detach(), "table": table}


def code_usage_stats(rep: torch.Tensor, codebook_size: int) -> dict:
    """Trainer statistics (larp_tokenizer_trainer.py:335-352; utils/statistics.py:5-32).
    ``rep`` (B, n) int64.  index_usage / kl_uni / perplexity use sample 0 only;
    index_usage_batch uses the whole local batch."""
    first = rep[0].flatten()
    counts = torch.bincount(first, minlength=codebook_size)
    counts_b = torch.bincount(rep.flatten(), minlength=codebook_size)
    usage = (counts > 0).float().sum().item() / codebook_size
    usage_b = (counts_b > 0).float().sum().item() / codebook_size
    total = counts.sum()
    if total == 0:
        kl = 0.0
    else:
        p = counts / total
        nz = p > 0
        kl = float((p[nz] * (torch.log(p[nz]) + math.log(codebook_size))).sum())
    avg = counts.float() / max(int(first.numel()), 1)
    perplexity = float((-(avg * torch.log(avg + 1e-10)).sum()).exp())
    return {"index_usage": usage, "index_usage_batch": usage_b, "kl_uni": kl,
            "perplexity": perplexity, "counts": counts, "counts_batch": counts_b}


def closed_form_backward(z, weight, idx, cfg: VQConfig, G, g_lossq: float):
    """Closed-form gradients of (regularized_z . G).sum() + g_lossq * loss_q with respect
    to raw z and the raw codebook; float64 so it can referee
    between autograd-on-the-restatement and the CUDA backward."""
    z64 = z.double().reshape(-1, z.shape[-1])
    W = weight.double()
    N, D = z64.shape
    G64 = G.double().reshape(N, D)
    if cfg.l2_normalized:
        zn = z64.norm(dim=1, keepdim=True)
        zc = zn.clamp_min(EPS)
        zh = z64 / zc
        wn = W.norm(dim=1, keepdim=True)
        wc = wn.clamp_min(EPS)
        E = W / wc
    else:
        zh, E = z64, W
    Q = E[idx.reshape(-1)]
    c = 2.0 * g_lossq / (N * D)
    dzh = G64 + cfg.commitment_loss_weight * c * (zh - Q)
    dE = torch.zeros_like(E)
    dE.index_add_(0, idx.reshape(-1), cfg.codebook_loss_weight * c * (Q - zh))
    if cfg.l2_normalized:
        proj = (zh * dzh).sum(1, keepdim=True)
        dz = torch.where(zn > EPS, (dzh - zh * proj) / zc, dzh / zc)
        projE = (E * dE).sum(1, keepdim=True)
        dW = torch.where(wn > EPS, (dE - E * projE) / wc, dE / wc)
    else:
        dz, dW = dzh, dE
    return dz.reshape(z.shape), dW

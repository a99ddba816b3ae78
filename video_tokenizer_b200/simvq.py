"""Drop-in for the reference's `SimVQ` quantizer (models/cosmos.py:772-876): a frozen
anchor table pushed through one learnable `Linear(e_dim, e_dim)`, nearest projected code by squared distance,
on the same sm_100a kernels as `SimpleVectorQuantizer`.

What this class restates on top of `functional.vq_quantize`:
  * the code table of a call is `embedding_proj(embedding.weight)` (cosmos.py:810) -- a (K, D) x (D, D) product
    done by torch; its gradient reaches `embedding_proj` through the `emb` / scatter-add outputs of the fused
    backward, the frozen `embedding.weight` gets none (cosmos.py:785);
  * ranking = squared distance on raw (un-normalised) vectors (cosmos.py:814-819) = MODE_ARGMIN_DIST with
    l2norm off, lowest index on ties like `torch.argmin`;
  * loss (cosmos.py:826-831): legacy    mean((q.detach()-z)^2) + beta * mean((q-z.detach())^2)
                              otherwise beta * mean((q.detach()-z)^2) + mean((q-z.detach())^2)
    = the kernels' loss_q with (commit weight, codebook weight) = (1, beta) or (beta, 1);
  * channel-first inputs (B,C,H,W) / (B,C,T,H,W) are moved to channel-last for the kernels and back
    (cosmos.py:795-803, 838-847); returns `(z_q, loss, (None, indices))` (cosmos.py:852-855).
The N x K distance matrix is never materialised.  Code dims up to b200vq_max_dim() = 1024 (the reference's
tokenizer builds this class with e_dim = 1024, cosmos.py:918-924) run on the any-D exact-fp32 kernels.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib
from . import functional as VF


class SimVQ(nn.Module):
    def __init__(self, n_e: int, e_dim: int, beta: float = 0.25, legacy: bool = True):
        super().__init__()
        self.n_e = n_e
        self.e_dim = e_dim
        self.beta = beta
        self.legacy = legacy
        self.embedding = nn.Embedding(self.n_e, self.e_dim)
        nn.init.normal_(self.embedding.weight, mean=0, std=self.e_dim ** -0.5)
        self.embedding.weight.requires_grad = False
        self.embedding_proj = nn.Linear(self.e_dim, self.e_dim)
        self.last_usage = None      # (K,) int32 device histogram of the last forward
        self.last_indices = None    # (N,) int64

    def _codebook(self):
        with torch.autocast(device_type="cuda", enabled=False):
            return F.linear(self.embedding.weight.float(), self.embedding_proj.weight.float(),
                            self.embedding_proj.bias.float())

    def forward(self, z):
        is_video = z.ndim == 5
        if z.ndim not in (4, 5):
            raise ValueError(f"SimVQ expects (B,C,H,W) or (B,C,T,H,W), got {tuple(z.shape)}")
        if z.shape[1] != self.e_dim:
            raise ValueError(f"channel dim {z.shape[1]} != e_dim {self.e_dim}")
        z_cl = (z.permute(0, 2, 3, 4, 1) if is_video else z.permute(0, 2, 3, 1)).contiguous()
        commit_w, codebook_w = (1.0, float(self.beta)) if self.legacy else (float(self.beta), 1.0)
        with torch.autocast(device_type="cuda", enabled=False):
            _zn, _emb, ste, idx, losses, hist, _hf = VF.vq_quantize(
                z_cl.reshape(1, -1, self.e_dim), self._codebook(), mode=_lib.MODE_ARGMIN_DIST, l2norm=False,
                beta=commit_w, w_cb=codebook_w, n_first=0)
        self.last_usage, self.last_indices = hist, idx
        z_q = ste.reshape(z_cl.shape)
        if is_video:
            z_q = z_q.permute(0, 4, 1, 2, 3).contiguous()
            indices = idx.view(z.shape[0], z.shape[2], z.shape[3], z.shape[4])
        else:
            z_q = z_q.permute(0, 3, 1, 2).contiguous()
            indices = idx.view(z.shape[0], z.shape[2], z.shape[3])
        return z_q, losses[0], (None, indices)

    def get_codebook_entry(self, indices):
        """indices (B,T,H,W) or (B,H,W) -> projected code vectors, channel-first (cosmos.py:857-876)"""
        table = self._codebook()
        if torch.is_grad_enabled() and table.requires_grad:
            z_q = F.embedding(indices, table)          # differentiable wrt embedding_proj, as the reference
        else:
            z_q = VF.codebook_lookup(indices, table.detach(), False)
        if z_q.ndim == 5:
            z_q = z_q.permute(0, 4, 1, 2, 3).contiguous()
        elif z_q.ndim == 4:
            z_q = z_q.permute(0, 3, 1, 2).contiguous()
        return z_q

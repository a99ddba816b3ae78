"""Drop-in for the reference's sibling quantizer `VectorQuantizer`
(models/model_new/quantizer/fsq.py:148-230; built as bottleneck_type 'sq' with the frozen K = 196 560
Leech-lattice table, D = 24, l2_norm=True, models/larp_tokenizer.py:230-233) on the same sm_100a
kernels as `SimpleVectorQuantizer`.

Differences from `SimpleVectorQuantizer` that this class restates on top of `functional.vq_quantize`:
  * loss = beta * mean(sum_d (q.detach() - z)^2) + mean(sum_d (q - z.detach())^2)   (fsq.py:207:
    summed over the code dimension before the mean) = D * (beta * L + L) with L the per-element mean
    the kernels return, so gradients are the closed form scaled by D;
  * returns only {'output' (straight-through, in the caller's layout), 'loss_codebook'} (fsq.py:212-213);
  * l2_norm=True ranks by -cos (fsq.py:186-189) = MODE_ARGMAX_COS, l2_norm=False by squared distance
    (fsq.py:191-192) = MODE_ARGMIN_DIST; lowest index on ties, as torch.argmin;
  * the usage histogram / entropy the reference computes and drops every call (fsq.py:195-196, 216-219)
    is a by-product of the fused finish kernel here: `last_usage` (K int32), `usage_entropy()` on demand.
The N x K distance matrix (196 560 x 4 B = 786 KB per token in the reference) is never materialised.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from . import functional as VF

_DEFAULT_TABLE = '/data2/zhxie/myproject/bsq-vit/cache/leech_lattices_normalized.npy'   # the reference's default (fsq.py:149)


class VectorQuantizer(nn.Module):
    def __init__(self, n_embed, embed_dim, l2_norm, beta, input_format='bchw', predefined_codebook=_DEFAULT_TABLE,
                 freeze_codebook=True):
        super().__init__()
        self.n_embed = n_embed
        self.embed_dim = embed_dim
        self.l2_norm = l2_norm
        self.beta = beta
        assert input_format in ['bchw', 'blc']
        self.input_format = input_format
        self.embedding = nn.Embedding(n_embed, embed_dim)
        self.embedding.weight.data.uniform_(-1 / n_embed, 1 / n_embed)
        self.bits_per_index = int(np.ceil(np.log2(n_embed)))
        if predefined_codebook is not None:
            table = torch.from_numpy(np.load(predefined_codebook))
            assert table.shape == (n_embed, embed_dim), 'Predefined codebook has incorrect shape'
            self.embedding.weight.data.copy_(table)
            self.embedding.weight.requires_grad = not freeze_codebook
        self.last_usage = None      # (K,) int32 device histogram of the last forward
        self.last_indices = None    # (N,) int64

    def forward(self, z):
        if self.input_format == 'bchw':
            z = z.permute(0, 2, 3, 1)
        if z.shape[-1] != self.embed_dim:
            raise ValueError(f"last dim {z.shape[-1]} != embed_dim {self.embed_dim}")
        mode = _lib.MODE_ARGMAX_COS if self.l2_norm else _lib.MODE_ARGMIN_DIST
        with torch.autocast(device_type="cuda", enabled=False):
            _zn, _emb, ste, idx, losses, hist, _hf = VF.vq_quantize(z, self.embedding.weight, mode=mode, l2norm=bool(self.l2_norm),
                                                                    beta=float(self.beta), w_cb=1.0, n_first=0)
        self.last_usage, self.last_indices = hist, idx
        loss = losses[0] * float(self.embed_dim)
        out = ste
        if self.input_format == 'bchw':
            out = out.permute(0, 3, 1, 2)
        return {'output': out, 'loss_codebook': loss}

    def usage_entropy(self, eps=1e-4):
        """get_entropy(bincount) of the last forward (fsq.py:195-196, 216-219), on device"""
        return self.get_entropy(self.last_usage.float(), eps)

    def get_entropy(self, count, eps=1e-4):
        probs = (count + eps) / (count + eps).sum()
        return -(probs * torch.log(probs)).sum()

    def get_codebook_entry(self, indices):
        z_q = VF.codebook_lookup(indices, self.embedding.weight, bool(self.l2_norm))
        if self.input_format == 'bchw':
            h = w = int(z_q.shape[1] ** 0.5)
            assert h * w == z_q.shape[1], 'Invalid sequence length'
            z_q = z_q.reshape(z_q.shape[0], h, w, z_q.shape[-1]).permute(0, 3, 1, 2)
        return z_q

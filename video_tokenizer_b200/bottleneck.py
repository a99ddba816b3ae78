"""B200-native drop-ins for the reference's ``models/bottleneck.py``.

Same class names, constructor arguments, attributes, state-dict keys and output
dictionaries as the reference (``Bottleneck`` :65-188, ``SimpleVectorQuantizer`` :203-344),
registered under the same registry names (``"bottleneck"``, ``"vq"``).  The arithmetic of
the quantizer runs in libb200vq.so (hand-written sm_100a CUDA); there is no CPU path.

Deliberate differences from the reference (documented in DESIGN.md):
  * ``input_norm_first`` / ``input_norm_last`` are 0-dim device tensors, not python
    floats, so ``Bottleneck.forward`` does no host sync (the trainer accepts either:
    trainers/larp_tokenizer_trainer.py:359-363).  ``sync_diagnostics=True`` restores floats.
  * stochastic + entropy_loss_weight > 0 raises UnboundLocalError up front, which is what
    the reference does at bottleneck.py:301.
  * ``regularizer.last_usage`` holds the fused code-usage histograms of the last forward.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib, functional as VF, registry
from .registry import register

_external_make = None  # set by registry.install(): the reference's own models.make


def entropy_loss(affinity, loss_type="softmax", temperature=0.01):
    """Reference entropy regulariser (bottleneck.py:12-33), kept in torch on the device: it
    is defined on the full N x K affinity matrix and its weight is 0.0 in every shipped cfg
    (SURVEY.md 8a-8).  Divides ``affinity`` in place like the reference."""
    flat = affinity.view(-1, affinity.shape[-1])
    flat /= temperature
    probs = F.softmax(flat, dim=-1)
    log_probs = F.log_softmax(flat + 1e-5, dim=-1)
    if loss_type == "softmax":
        target_probs = probs
    elif loss_type == "argmax":
        codes = torch.argmax(flat, dim=-1)
        onehots = F.one_hot(codes, num_classes=flat.shape[-1]).to(flat.dtype)
        target_probs = probs - (probs - onehots).detach()
    else:
        raise ValueError(f"Entropy loss {loss_type} not supported")
    avg_probs = torch.mean(target_probs, dim=0)
    avg_entropy = -torch.sum(avg_probs * torch.log(avg_probs + 1e-5))
    sample_entropy = -torch.mean(torch.sum(target_probs * log_probs, dim=-1))
    return sample_entropy - avg_entropy, sample_entropy, avg_entropy


class AbstractRegularizer(nn.Module):
    def __init__(self, *args, **kwargs):
        super().__init__()

    def forward(self, z):
        raise NotImplementedError

    def decode(self, z):
        raise NotImplementedError


_ZEROS: dict = {}


def _zero_scalar(device):
    """0-dim fp32 zero for the three entropy keys when entropy_loss_weight == 0 (bottleneck.py:303 builds a new
    tensor per call; a fill kernel + ~10 us of host time per step here).  Shared and read-only by convention."""
    z = _ZEROS.get(device)
    if z is None:
        z = _ZEROS[device] = torch.zeros((), device=device, dtype=torch.float32)
    return z


@register("vq")
class SimpleVectorQuantizer(AbstractRegularizer):
    """Drop-in for the reference quantizer (bottleneck.py:203-344)."""

    def __init__(self, dim, codebook_size, commitment_loss_weight=0.25, entropy_loss_weight=0.0,
                 entropy_loss_temperature=0.01, l2_normalized=False, same_index_shape=True, stochastic=False,
                 stochastic_temperature=1.0, codebook_loss_weight=1.0, **kwargs):
        super().__init__()
        self.codebook_size = codebook_size
        self.dim = dim
        self.beta = commitment_loss_weight
        self.codebook_loss_weight = codebook_loss_weight
        self.entropy_loss_weight = entropy_loss_weight
        assert isinstance(l2_normalized, bool)
        self.l2_normalized = l2_normalized
        self.stochastic = stochastic
        self.eval_deterministic = False
        self.default_stochastic_temperature = stochastic_temperature
        if self.stochastic:
            if stochastic_temperature > 0:  # fixed temperature
                self.stochastic_temperature_inv = 1 / stochastic_temperature
            else:  # learnable temperature (receives no gradient: sampling is not differentiable)
                self.stochastic_temperature_inv = nn.Parameter(torch.tensor(10.0))
        self.embedding = nn.Embedding(self.codebook_size, self.dim)
        nn.init.kaiming_uniform_(self.embedding.weight)
        self.same_index_shape = same_index_shape
        self.entropy_loss_temperature = entropy_loss_temperature
        # B200 extras (accepted through **kwargs by the reference signature too)
        self.return_usage = bool(kwargs.get("return_usage", False))
        self.last_usage = None  # (hist_first, hist_batch) int32 (K,) of the last forward
        self.philox_dev = None  # int64[2] device {seed, offset} when the RNG state lives on the GPU (CUDA graphs)
        if dim > 1024:
            raise ValueError(f"bottleneck dim {dim} > 1024 is not supported by the B200 quantizer")

    def set_eval_deterministic(self, deterministic=True):
        self.eval_deterministic = deterministic

    def set_stochastic_temperature(self, temperature):
        self.stochastic_temperature_inv = 1 / temperature

    @torch.autocast(device_type="cuda", enabled=False)
    def get_emb(self):
        w = self.embedding.weight
        assert w.dtype == torch.float32, f"Embedding weight dtype is {w.dtype}, expected float32"
        if self.l2_normalized:
            # differentiable (callers of get_emb may backprop); forward() uses the fused kernel
            return F.normalize(w, p=2, dim=-1)
        return w

    def enable_device_rng(self, device=None):
        """Move the sampler's Philox (seed, offset) to the GPU: the kernels read it there and advance the offset
        themselves, so a captured CUDA graph draws fresh noise on every replay.  The state starts from the torch
        CUDA generator (and consumes nothing from it afterwards); call again to re-synchronise with the generator.
        Only the one-call `Bottleneck` path on the tensor-core sampler reads it."""
        dev = torch.device(device) if device is not None else self.embedding.weight.device
        gen = torch.cuda.default_generators[dev.index if dev.index is not None else torch.cuda.current_device()]
        seed = gen.initial_seed() & 0xFFFFFFFFFFFFFFFF
        state = torch.tensor([seed if seed < 2 ** 63 else seed - 2 ** 64, gen.get_offset()], dtype=torch.int64)
        if self.philox_dev is None or self.philox_dev.device != dev:
            self.philox_dev = state.to(dev)
        else:
            self.philox_dev.copy_(state)
        return self.philox_dev

    def disable_device_rng(self):
        self.philox_dev = None

    def _mode(self):
        if self.stochastic:
            return _lib.MODE_ARGMAX_COS if (self.eval_deterministic and not self.training) else _lib.MODE_SAMPLE
        return _lib.MODE_ARGMIN_DIST

    def _step_config(self, device, device_rng_ok=False):
        """(mode, 1/T, philox seed, philox offset) of this call, after the reference's checks.  With
        `device_rng_ok` (the one-call Bottleneck path) a GPU-resident state (`enable_device_rng`) replaces the
        host-side draw from the torch generator; every other path keeps using the generator."""
        if self.stochastic:
            assert self.l2_normalized, "Stochastic sampling requires l2 normalization"
            if self.entropy_loss_weight > 0:
                raise UnboundLocalError("cannot access local variable 'd' where it is not associated with a value")
        mode = self._mode()
        t_inv = float(self.stochastic_temperature_inv) if self.stochastic else 1.0
        seed = offset = 0
        if mode == _lib.MODE_SAMPLE and not (device_rng_ok and self.philox_dev is not None):
            seed, offset = VF.philox_state(device)
        return mode, t_inv, seed, offset

    def _assemble(self, z_shape, znorm, emb, ste, idx, losses, hist, hist_first, usage=None):
        """Output dict of forward() (bottleneck.py:312-323) for entropy_loss_weight == 0."""
        self.last_usage = (hist_first, hist)
        self._last_stats = usage
        zero = _zero_scalar(znorm.device)
        out = {
            "unregularized_z": znorm,
            "emb": emb,
            "regularized_z": ste,
            "bottleneck_rep": idx.reshape(z_shape[0], z_shape[1]) if self.same_index_shape else idx,
            "loss_q": losses[0],
            "loss_commit": losses[1],
            "loss_codebook": losses[2],
            "loss_entropy": zero,
            "per_sample_entropy": zero,
            "codebook_entropy": zero,
        }
        if self.return_usage:
            st = (usage if usage is not None else VF.usage_stats(hist_first, hist)).unbind(0)   # one op, four views
            out.update(index_usage=st[0], index_usage_batch=st[1], kl_uni=st[2], perplexity=st[3])
        return out

    @torch.autocast(device_type="cuda", enabled=False)
    def forward(self, z):
        assert len(z.shape) == 3, "Input shape must be (batch, n_tokens, e_dim)"
        VF._require_cuda(z, "z")
        mode, t_inv, seed, offset = self._step_config(z.device)
        znorm, emb, ste, idx, losses, hist, hist_first = VF.vq_quantize(
            z, self.embedding.weight, mode=mode, l2norm=self.l2_normalized, t_inv=t_inv, beta=self.beta,
            w_cb=self.codebook_loss_weight, n_first=z.shape[1], seed=seed, offset=offset)
        self.last_usage = (hist_first, hist)
        self._last_stats = None
        loss_commit, loss_codebook = losses[1], losses[2]
        if self.entropy_loss_weight > 0:
            if VF.entropy_supported(self.dim):
                # streamed: the N x K affinity / softmax matrices of reference :300-303 never exist
                loss_entropy, sample_entropy, avg_entropy = VF.entropy_loss_streamed(
                    znorm, emb, self.entropy_loss_temperature)
            else:
                # code dims without a streaming specialisation: the reference's N x K formulation in torch
                zf = znorm.reshape(-1, self.dim)
                d = (torch.sum(zf ** 2, dim=1, keepdim=True) + torch.sum(emb ** 2, dim=1)
                     - 2 * torch.einsum("bd,dn->bn", zf, emb.t()))
                loss_entropy, sample_entropy, avg_entropy = entropy_loss(-d, temperature=self.entropy_loss_temperature)
            loss = self.beta * loss_commit + self.codebook_loss_weight * loss_codebook \
                + self.entropy_loss_weight * loss_entropy
        else:
            loss_entropy = sample_entropy = avg_entropy = torch.tensor(0.0, device=z.device, dtype=torch.float32)
            loss = losses[0]
        q_indices = idx.reshape(z.shape[0], z.shape[1]) if self.same_index_shape else idx
        out = {
            "unregularized_z": znorm,  # l2 normalized if l2_normalized=True
            "emb": emb,
            "regularized_z": ste,
            "bottleneck_rep": q_indices,
            "loss_q": loss,
            "loss_commit": loss_commit,
            "loss_codebook": loss_codebook,
            "loss_entropy": loss_entropy,
            "per_sample_entropy": sample_entropy,
            "codebook_entropy": avg_entropy,
        }
        if self.return_usage:
            st = VF.usage_stats(hist_first, hist)
            out.update(index_usage=st[0], index_usage_batch=st[1], kl_uni=st[2], perplexity=st[3])
        return out

    def code_usage(self):
        """Device tensor (4,): index_usage, index_usage_batch, kl_uni, perplexity of the last
        forward (trainers/larp_tokenizer_trainer.py:335-352), computed from the fused histograms."""
        if self.last_usage is None:
            raise RuntimeError("code_usage() needs a forward pass first")
        if getattr(self, "_last_stats", None) is not None:   # the one-call Bottleneck step already made them
            return self._last_stats
        return VF.usage_stats(*self.last_usage)

    def get_codebook_entry(self, indices, shape=None):
        """indices -> (normalised) code rows (bottleneck.py:327-341).  Inference goes through b200vq_decode; when a
        gradient towards the codebook is being recorded the lookup is the reference's differentiable op sequence
        (nn.Embedding + F.normalize), because the decode kernel builds no autograd graph."""
        w = self.embedding.weight
        if torch.is_grad_enabled() and w.requires_grad:
            z_q = self.embedding(indices.reshape(-1))
            if self.l2_normalized:
                z_q = F.normalize(z_q, p=2, dim=-1)
            return z_q.reshape(shape) if shape is not None else z_q.reshape(*indices.shape, self.dim)
        return VF.codebook_lookup(indices, w, self.l2_normalized, shape)

    def decode(self, indices):
        return self.get_codebook_entry(indices)


@register("bottleneck")
class Bottleneck(nn.Module):
    """Drop-in for the reference Bottleneck (bottleneck.py:65-188)."""

    def __init__(self, bottleneck_dim: int, input_dim: int, output_dim: int, token_nums: int, norm=None,
                 regularizer=None, param_names=None, sync_diagnostics: bool = False):
        super().__init__()
        self.token_nums = token_nums
        self.param_names = param_names
        self.input_dim = input_dim
        self.output_dim = output_dim
        self.sync_diagnostics = sync_diagnostics
        if bottleneck_dim > 0:
            self.bottleneck_dim = bottleneck_dim
        else:
            assert self.input_dim == self.output_dim, \
                "input_dim and output_dim must be the same when bottleneck_dim is not specified"
            self.bottleneck_dim = self.input_dim
        if norm is not None:
            norm = norm.lower()
            self.norm = None if norm in ("no", "none") else norm
        else:
            self.norm = None
        if regularizer is not None and "kl" in regularizer["name"].lower() \
                and regularizer["name"].lower() not in ["vqkl"]:
            self.project_dim = self.bottleneck_dim * 2
        else:
            self.project_dim = self.bottleneck_dim
        assert isinstance(self.input_dim, int) and isinstance(self.token_nums, int), \
            "input_dim and n_tokens must be integers"
        self.param_bottnecks = None
        self.in_linear = nn.Linear(self.input_dim, self.project_dim)
        self.out_linear = nn.Linear(self.bottleneck_dim, self.output_dim)
        if self.norm is None:
            self.norm_layer = lambda x: x
        elif self.norm == "bn_bn":
            self.norm_layer = nn.SyncBatchNorm(self.project_dim)
        elif self.norm == "bn_b":
            assert self.token_nums is not None, "num_tokens must be specified for batch normalization"
            self.norm_layer = nn.SyncBatchNorm(self.project_dim * self.token_nums)
        elif self.norm == "ln_d":
            self.norm_layer = nn.LayerNorm(self.project_dim)
        elif self.norm == "ln_nd":
            self.norm_layer = nn.LayerNorm((self.token_nums, self.project_dim))
        elif self.norm == "ln_d_na":
            self.norm_layer = nn.LayerNorm(self.project_dim, elementwise_affine=False)
        else:
            raise ValueError(f"Normalization type {norm} not supported")
        if regularizer is not None and regularizer["name"].lower() not in ["no", "none"]:
            regularizer["args"]["dim"] = self.bottleneck_dim
            regularizer["args"]["token_nums"] = self.token_nums
            mk = _external_make or registry.make
            self.regularizer = mk(regularizer)
        else:
            self.regularizer = None

    def project_in(self, x):
        assert len(x.shape) == 3, "Input shape must be (batch, n_tokens, e_dim)"
        z = VF.linear_in(x, self.in_linear.weight, self.in_linear.bias)
        if self.norm is not None:
            with torch.autocast(device_type="cuda", enabled=False):
                z = z.float()
                if self.norm == "bn_bn":
                    z = self.norm_layer(z.transpose(1, 2)).transpose(1, 2)
                elif self.norm == "bn_b":
                    b, n, d = z.shape
                    z = self.norm_layer(z.reshape(b, n * d)).reshape(b, n, d)
                elif self.norm in ("ln_nd", "ln_d"):
                    z = self.norm_layer(z)
                # 'ln_d_na' is constructed but never applied by the reference (:143-160)
        return z

    def project_out(self, z_cat):
        return VF.linear_out(z_cat, self.out_linear.weight, self.out_linear.bias)

    def decode(self, bottleneck_rep, add=None):
        """indices -> hidden (bottleneck.py:166-168).  `add` (optional, (token_nums, C)): folded into the
        store of the fused kernel -- the tokenizer's `z + decoder_latent_pe` (larp_tokenizer.py:463-464)."""
        reg = self.regularizer
        w = self.out_linear.weight
        if (type(reg) is SimpleVectorQuantizer and bottleneck_rep.is_cuda and not bottleneck_rep.dtype.is_floating_point
                and w.dtype == torch.float32 and reg.embedding.weight.dtype == torch.float32
                and VF.linear_dims_supported(w.shape[0], w.shape[1]) and w.shape[1] == reg.dim
                and not (torch.is_grad_enabled() and (w.requires_grad or reg.embedding.weight.requires_grad))):
            return VF.decode_project(bottleneck_rep, reg.embedding.weight, reg.l2_normalized, w, self.out_linear.bias, add)
        regularized_z = self.regularizer.decode(bottleneck_rep)
        out = self.project_out(regularized_z)
        return out if add is None else out + add

    def _input_norms(self, x):
        with torch.no_grad():
            first = torch.linalg.vector_norm(x[:, 0, :].float(), dim=-1).mean()
            last = torch.linalg.vector_norm(x[:, -1, :].float(), dim=-1).mean()
        return (first.item(), last.item()) if self.sync_diagnostics else (first, last)

    def forward(self, x, add=None):
        """Reference Bottleneck.forward (bottleneck.py:170-188).  `add` (optional, B200 extra): an (n, C) table added to
        'output' inside the out_linear store -- the tokenizer's `z + decoder_latent_pe` (models/larp_tokenizer.py:463-464)
        for callers that want that pass folded in; treated as a constant (it is a buffer in every shipped config)."""
        reg = self.regularizer
        if (self.norm is None and type(reg) is SimpleVectorQuantizer and reg.entropy_loss_weight == 0
                and self.project_dim == self.bottleneck_dim and len(x.shape) == 3 and self.in_linear.bias is not None
                and VF.bottleneck_supported(x, self.in_linear.weight, reg.embedding.weight, self.out_linear.weight)):
            # one host call per direction: in_linear -> quantizer -> out_linear (+ usage statistics)
            mode, t_inv, seed, offset = reg._step_config(x.device, device_rng_ok=True)
            x_hat, projected_z, znorm, emb, ste, idx, lq, lc, lcb, hists, usage, norms = VF.bottleneck_step(
                x, self.in_linear.weight, self.in_linear.bias, reg.embedding.weight, self.out_linear.weight,
                self.out_linear.bias, mode=mode, l2norm=reg.l2_normalized, t_inv=t_inv, beta=reg.beta,
                w_cb=reg.codebook_loss_weight, n_first=x.shape[1], seed=seed, offset=offset,
                philox_dev=reg.philox_dev if mode == _lib.MODE_SAMPLE else None, add=add)
            regularized_output = reg._assemble(projected_z.shape, znorm, emb, ste, idx, (lq, lc, lcb), hists[0], hists[1], usage)
            bottleneck_rep = regularized_output.pop("bottleneck_rep")
            input_norm_first, input_norm_last = norms[0], norms[1]
            if self.sync_diagnostics:
                input_norm_first, input_norm_last = input_norm_first.item(), input_norm_last.item()
            return {
                "output": x_hat,
                "bottleneck_rep": bottleneck_rep,
                "projected_z": projected_z,
                "input_norm_first": input_norm_first,
                "input_norm_last": input_norm_last,
                **regularized_output,
            }
        input_norm_first, input_norm_last = self._input_norms(x)
        z = self.project_in(x)
        projected_z = z
        if self.regularizer is not None:
            regularized_output = self.regularizer(z)
        else:
            regularized_output = {"regularized_z": z}
        x_hat = self.project_out(regularized_output["regularized_z"])
        if add is not None:
            x_hat = x_hat + add.detach()
        bottleneck_rep = regularized_output.pop("bottleneck_rep")  # KeyError without a regularizer, as in the reference
        return {
            "output": x_hat,
            "bottleneck_rep": bottleneck_rep,
            "projected_z": projected_z,
            "input_norm_first": input_norm_first,
            "input_norm_last": input_norm_last,
            **regularized_output,
        }

"""Autograd-level wrappers over the C ABI (include/b200vq.h).

Torch is plumbing here: it owns device memory, streams and the autograd tape; every
numeric step of the quantizer runs in libb200vq.so.  There is no CPU path: tensors that
are not on a CUDA device raise.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from . import _lib

_DTYPES = {torch.float32: _lib.DT_F32, torch.bfloat16: _lib.DT_BF16, torch.float16: _lib.DT_F16}
_workspaces: dict = {}
# device index -> torch.cuda.Event the one-call backward records when its parameter gradients are enqueued
# (sharded.enable_grad_overlap); absent = no event
GRADS_READY: dict = {}

# "tc": tcgen05 tensor-core scoring where supported (default); "exact": fp32 CUDA-core scoring.
# Deterministic indices are identical either way (tests/test_gpu_tc.py).
SCORING = os.environ.get("B200VQ_SCORING", "tc")


def set_scoring(kind: str) -> str:
    """Select the scoring kernels: "tc" (tensor cores, default) or "exact" (fp32 CUDA cores)."""
    global SCORING
    if kind not in ("tc", "exact"):
        raise ValueError("scoring must be 'tc' or 'exact'")
    prev, SCORING = SCORING, kind
    return prev


def _require_cuda(t: torch.Tensor, name: str):
    if not t.is_cuda:
        raise RuntimeError(f"video_tokenizer_b200: '{name}' must be a CUDA tensor (the B200 path has no CPU fallback)")


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream(index: Optional[int] = None) -> int:
    """cudaStream_t of torch's current stream on the current (or given) device.  The raw getter avoids building a
    torch.cuda.Stream object per call (the host side of a step is ~0.5 ms: every 10 us counts)."""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device() if index is None else index)
    return torch.cuda.current_stream(index).cuda_stream


class _NullGuard:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NULL_GUARD = _NullGuard()


def _on_device(dev: torch.device):
    """`with torch.cuda.device(dev)` only when dev is not already current (the guard costs ~15 us per use)"""
    return _NULL_GUARD if torch.cuda.current_device() == dev.index else torch.cuda.device(dev)


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    key = (device.index, _stream(device.index))
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def rows_view(z: torch.Tensor) -> Tuple[torch.Tensor, int]:
    """(tensor, row stride in elements) such that row i of the flattened (N, D) view starts
    at data_ptr + i*stride.  Copies only when the layout is not row-uniform."""
    D = z.shape[-1]
    if z.dim() == 2:
        if z.stride(1) == 1 and z.stride(0) >= D:
            return z, z.stride(0)
    elif z.dim() == 3:
        B, n, _ = z.shape
        if z.stride(2) == 1 and z.stride(1) >= D and (B == 1 or z.stride(0) == n * z.stride(1)):
            return z, z.stride(1)
    zc = z.contiguous()
    return zc, D


def philox_state(device: torch.device) -> Tuple[int, int]:
    """(seed, offset) from the torch CUDA generator, advancing it, so checkpointed RNG
    state reproduces the stochastic draws (reference consumes the same generator through
    torch.multinomial, bottleneck.py:280)."""
    gen = torch.cuda.default_generators[device.index if device.index is not None else torch.cuda.current_device()]
    seed = gen.initial_seed() & 0xFFFFFFFFFFFFFFFF
    offset = gen.get_offset()
    gen.set_offset(offset + 4)
    return seed, offset


def prepare_codebook(weight: torch.Tensor, l2norm: bool):
    """-> emb (K,D), e2 (K), wnorm (K)  [get_emb, bottleneck.py:252-259]"""
    _require_cuda(weight, "embedding.weight")
    assert weight.dtype == torch.float32, f"Embedding weight dtype is {weight.dtype}, expected float32"
    w = weight.detach().contiguous()
    K, D = w.shape
    emb = torch.empty_like(w)
    e2 = torch.empty(K, dtype=torch.float32, device=w.device)
    wnorm = torch.empty(K, dtype=torch.float32, device=w.device)
    lib = _lib.load()
    with _on_device(w.device):
        _lib.check(lib.b200vq_prepare_codebook(w.data_ptr(), K, D, int(l2norm), emb.data_ptr(), e2.data_ptr(),
                                               wnorm.data_ptr(), _stream()), "b200vq_prepare_codebook")
    return emb, e2, wnorm


def prepare_codebook_tc(emb: torch.Tensor) -> Optional[torch.Tensor]:
    """Split-bf16 tile image of the normalised codebook for the tcgen05 path, or None when the
    code dimension has no tensor-core specialisation."""
    K, D = emb.shape
    lib = _lib.load()
    nbytes = lib.b200vq_tc_image_bytes(K, D)
    if nbytes == 0:
        return None
    image = torch.empty(nbytes, dtype=torch.uint8, device=emb.device)
    with _on_device(emb.device):
        _lib.check(lib.b200vq_prepare_codebook_tc(emb.data_ptr(), K, D, image.data_ptr(), _stream()),
                   "b200vq_prepare_codebook_tc")
    return image


class _VQFunction(torch.autograd.Function):
    """forward = b200vq_prepare_codebook + b200vq_forward; backward = b200vq_backward +
    b200vq_codebook_backward.  Outputs: znorm, emb, ste, idx, losses(3), hist, hist_first."""

    @staticmethod
    def forward(ctx, z, weight, mode, l2norm, t_inv, beta, w_cb, n_first, seed, offset):
        _require_cuda(z, "z")
        _require_cuda(weight, "embedding.weight")
        if z.dtype not in _DTYPES:
            raise ValueError(f"unsupported latent dtype {z.dtype}")
        lib = _lib.load()
        dev = z.device
        D = z.shape[-1]
        K = weight.shape[0]
        if weight.shape[1] != D:
            raise ValueError(f"latent dim {D} != codebook dim {weight.shape[1]}")
        zz, stride = rows_view(z.detach())
        N = zz.shape[:-1].numel()
        with _on_device(dev):
            emb, e2, wnorm = prepare_codebook(weight, l2norm)
            image = None
            if SCORING == "tc" and lib.b200vq_tc_supported(D, int(l2norm), mode):
                image = prepare_codebook_tc(emb)
            znorm = torch.empty(*z.shape, dtype=torch.float32, device=dev)
            ste = torch.empty(*z.shape, dtype=torch.float32, device=dev)
            idx = torch.empty(N, dtype=torch.int64, device=dev)
            zlen = torch.empty(N, dtype=torch.float32, device=dev)
            hist = torch.empty(K, dtype=torch.int32, device=dev)
            hist_first = torch.empty(K, dtype=torch.int32, device=dev)
            losses = torch.empty(3, dtype=torch.float32, device=dev)
            nbytes = lib.b200vq_workspace_bytes(N, D, K)
            ws = _workspace(dev, nbytes)
            rc = lib.b200vq_forward(zz.data_ptr(), _DTYPES[zz.dtype], stride, N, D, emb.data_ptr(), e2.data_ptr(),
                                    _ptr(image), K, mode, int(l2norm), float(t_inv), seed, offset, int(n_first), float(beta),
                                    float(w_cb), znorm.data_ptr(), ste.data_ptr(), idx.data_ptr(), zlen.data_ptr(),
                                    hist.data_ptr(), hist_first.data_ptr(), losses.data_ptr(), ws.data_ptr(),
                                    ws.numel(), _stream())
            _lib.check(rc, "b200vq_forward")
        ctx.save_for_backward(znorm, zlen, idx, emb, wnorm)
        ctx.cfg = (bool(l2norm), float(beta), float(w_cb), N, D, K, z.dtype, tuple(z.shape))
        ctx.mark_non_differentiable(idx, hist, hist_first)
        return znorm, emb, ste, idx, losses, hist, hist_first

    @staticmethod
    def backward(ctx, g_znorm, g_emb, g_ste, _g_idx, g_losses, _g_h, _g_hf):
        znorm, zlen, idx, emb, wnorm = ctx.saved_tensors
        l2norm, beta, w_cb, N, D, K, z_dtype, z_shape = ctx.cfg
        lib = _lib.load()
        dev = znorm.device

        def _f32c(t):
            return None if t is None else t.to(torch.float32).contiguous()

        g_znorm, g_emb, g_ste, g_losses = _f32c(g_znorm), _f32c(g_emb), _f32c(g_ste), _f32c(g_losses)
        need_z, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        grad_z = grad_w = None
        with _on_device(dev):
            gz = torch.empty(z_shape, dtype=z_dtype, device=dev)
            grad_emb = torch.empty(K, D, dtype=torch.float32, device=dev)
            rc = lib.b200vq_backward(_ptr(g_ste), _ptr(g_znorm), _ptr(g_losses), znorm.data_ptr(), zlen.data_ptr(),
                                     idx.data_ptr(), emb.data_ptr(), None, 0, N, D, K, int(l2norm), N, beta, w_cb,
                                     gz.data_ptr(), _DTYPES[z_dtype], grad_emb.data_ptr(), _stream())
            _lib.check(rc, "b200vq_backward")
            if need_z:
                grad_z = gz
            if need_w:
                grad_w = torch.empty(K, D, dtype=torch.float32, device=dev)
                rc = lib.b200vq_codebook_backward(grad_emb.data_ptr(), _ptr(g_emb), emb.data_ptr(), wnorm.data_ptr(),
                                                  K, D, int(l2norm), grad_w.data_ptr(), _stream())
                _lib.check(rc, "b200vq_codebook_backward")
        return grad_z, grad_w, None, None, None, None, None, None, None, None


def vq_quantize(z, weight, *, mode, l2norm, t_inv=1.0, beta=0.25, w_cb=1.0, n_first=None, seed=0, offset=0):
    """Fused quantizer step.  Returns (znorm, emb, ste, idx(N,), losses(3)=[loss_q,
    loss_commit, loss_codebook], hist(K) int32, hist_first(K) int32)."""
    if n_first is None:
        n_first = z.shape[1] if z.dim() == 3 else z.shape[0]
    return _VQFunction.apply(z, weight, int(mode), bool(l2norm), float(t_inv), float(beta), float(w_cb),
                             int(n_first), int(seed), int(offset))


class _EntropyFunction(torch.autograd.Function):
    """Streaming `entropy_loss(-d, temperature=T_e)` (reference models/bottleneck.py:12-33 as called at
    :300-303): b200vq_entropy_forward / _backward.  Inputs are the tensors the reference builds d from
    ('unregularized_z' rows and 'emb'); the N x K affinity / softmax / log-softmax matrices never exist."""

    @staticmethod
    def forward(ctx, znorm, emb, t_inv):
        _require_cuda(znorm, "z")
        _require_cuda(emb, "emb")
        lib = _lib.load()
        dev = znorm.device
        D = znorm.shape[-1]
        K = emb.shape[0]
        zz = znorm.detach().reshape(-1, D).to(torch.float32).contiguous()
        ee = emb.detach().to(torch.float32).contiguous()
        N = zz.shape[0]
        f32 = dict(dtype=torch.float32, device=dev)
        with _on_device(dev):
            lse = torch.empty(N, **f32)
            ent = torch.empty(N, **f32)
            psum = torch.empty(K, **f32)
            w = torch.empty(K, **f32)
            ent_sum = torch.empty(1, dtype=torch.float64, device=dev)
            out3 = torch.empty(3, **f32)
            nscr = lib.b200vq_entropy_scratch_bytes(N, K)
            scr = torch.empty(max(nscr, 16), dtype=torch.uint8, device=dev)
            rc = lib.b200vq_entropy_forward(zz.data_ptr(), N, D, ee.data_ptr(), K, float(t_inv), lse.data_ptr(),
                                            ent.data_ptr(), psum.data_ptr(), w.data_ptr(), ent_sum.data_ptr(),
                                            out3.data_ptr(), scr.data_ptr(), nscr, _stream())
            _lib.check(rc, "b200vq_entropy_forward")
        ctx.save_for_backward(zz, ee, lse, ent, w)
        ctx.cfg = (float(t_inv), N, D, K, tuple(znorm.shape), tuple(emb.shape))
        l, s, a = out3.unbind(0)
        return l, s, a

    @staticmethod
    def backward(ctx, g_loss, g_sample, g_avg):
        zz, ee, lse, ent, w = ctx.saved_tensors
        t_inv, N, D, K, z_shape, e_shape = ctx.cfg
        lib = _lib.load()
        dev = zz.device
        f32 = dict(dtype=torch.float32, device=dev)

        def _s(t):
            return None if t is None else t.to(torch.float32).contiguous()

        g_loss, g_sample, g_avg = _s(g_loss), _s(g_sample), _s(g_avg)
        need_z, need_e = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        gz = ge = None
        with _on_device(dev):
            g3 = torch.empty(3, **f32)
            _lib.check(lib.b200vq_pack3(_ptr(g_loss), _ptr(g_sample), _ptr(g_avg), g3.data_ptr(), _stream()), "b200vq_pack3")
            c = torch.empty(N, **f32)
            if need_z:
                gz = torch.empty(N, D, **f32)
            if need_e:
                ge = torch.empty(K, D, **f32)
            rc = lib.b200vq_entropy_backward(zz.data_ptr(), N, D, ee.data_ptr(), K, t_inv, lse.data_ptr(), ent.data_ptr(),
                                             w.data_ptr(), g3.data_ptr(), c.data_ptr(), _ptr(gz), _ptr(ge), _stream())
            _lib.check(rc, "b200vq_entropy_backward")
        return (gz.reshape(z_shape) if gz is not None else None), (ge.reshape(e_shape) if ge is not None else None), None


def entropy_supported(D: int) -> bool:
    return bool(_lib.load().b200vq_entropy_supported(int(D)))


def entropy_loss_streamed(znorm, emb, temperature=0.01):
    """(loss_entropy, per_sample_entropy, codebook_entropy) of reference `entropy_loss(-d, temperature)`
    (models/bottleneck.py:12-33, :300-303) with d the squared distances between `znorm` rows and `emb` rows."""
    return _EntropyFunction.apply(znorm, emb, 1.0 / float(temperature))


# "fused": the Bottleneck's in/out linears run through the streaming kernels of libb200vq.so when
# the shapes allow (fp32, D in {8,16,24,32}, C % 16 == 0, no autocast); "torch": always F.linear.
LINEARS = os.environ.get("B200VQ_LINEARS", "fused")


def set_linears(kind: str) -> str:
    global LINEARS
    if kind not in ("fused", "torch"):
        raise ValueError("linears must be 'fused' or 'torch'")
    prev, LINEARS = LINEARS, kind
    return prev


def linears_supported(x: torch.Tensor, weight: torch.Tensor, C: int, D: int) -> bool:
    if LINEARS != "fused" or not x.is_cuda or x.dtype != torch.float32 or weight.dtype != torch.float32:
        return False
    if torch.is_autocast_enabled():
        return False
    return bool(_lib.load().b200vq_linear_supported(C, D))


def linear_dims_supported(C: int, D: int) -> bool:
    return LINEARS == "fused" and not torch.is_autocast_enabled() and bool(_lib.load().b200vq_linear_supported(C, D))


def _rows2d(x: torch.Tensor, gran: Optional[int] = None) -> Tuple[torch.Tensor, int, int]:
    """(tensor, row stride, N) for a (..., C) tensor viewed as N rows; copies unless row-uniform
    with a 16-byte aligned base and a stride that is a multiple of `gran` elements (16 bytes' worth)."""
    C = x.shape[-1]
    gran = gran or max(1, 16 // x.element_size())
    t, stride = rows_view(x) if x.dim() in (2, 3) else (x.contiguous(), C)
    if stride % gran or t.data_ptr() % 16:
        t, stride = x.contiguous(), C
    return t, stride, t.shape[:-1].numel()


def _wg_workspace(lib, C, D, device):
    """slabs of per-CTA partial weight gradients (b200vq_weight_grad_ws: no atomics, deterministic)"""
    return torch.empty(int(lib.b200vq_weight_grad_workspace_bytes(C, D)), dtype=torch.uint8, device=device)


class _LinearIn(torch.autograd.Function):
    """z = x W^T + b with W (D, C): b200vq_linear_in; backward = b200vq_linear_out (grad_x),
    b200vq_weight_grad (grad_W as (D, C)), b200vq_colsum (grad_b)."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        lib = _lib.load()
        D, C = weight.shape
        xx, ldx, N = _rows2d(x.detach())
        w = weight.detach().contiguous()
        y = torch.empty(*x.shape[:-1], D, dtype=torch.float32, device=x.device)
        with _on_device(x.device):
            _lib.check(lib.b200vq_linear_in(xx.data_ptr(), ldx, N, C, w.data_ptr(), _ptr(bias.detach() if bias is not None else None),
                                            D, y.data_ptr(), _stream()), "b200vq_linear_in")
        ctx.save_for_backward(x, weight)
        ctx.has_bias = bias is not None
        return y

    @staticmethod
    def backward(ctx, gy):
        x, weight = ctx.saved_tensors
        lib = _lib.load()
        D, C = weight.shape
        gy = gy.to(torch.float32).contiguous()
        N = gy.numel() // D
        gx = gw = gb = None
        with _on_device(gy.device):
            w = weight.detach().contiguous()
            if ctx.needs_input_grad[0]:
                gx = torch.empty(x.shape, dtype=torch.float32, device=gy.device)
                _lib.check(lib.b200vq_linear_out(gy.data_ptr(), N, D, w.data_ptr(), 1, None, C, gx.data_ptr(), C, _stream()),
                           "b200vq_linear_out(grad_x)")
            if ctx.needs_input_grad[1]:
                xx, ldx, _ = _rows2d(x.detach())
                gw = torch.empty(D, C, dtype=torch.float32, device=gy.device)
                ws = _wg_workspace(lib, C, D, gy.device)
                _lib.check(lib.b200vq_weight_grad_ws(xx.data_ptr(), ldx, gy.data_ptr(), N, C, D, gw.data_ptr(), 1, None,
                                                     ws.data_ptr(), ws.numel(), _stream()), "b200vq_weight_grad_ws(in_linear)")
            if ctx.has_bias and ctx.needs_input_grad[2]:
                gb = torch.empty(D, dtype=torch.float32, device=gy.device)
                _lib.check(lib.b200vq_colsum(gy.data_ptr(), N, D, gb.data_ptr(), _stream()), "b200vq_colsum")
        return gx, gw, gb


class _LinearOut(torch.autograd.Function):
    """y = a W^T + b with W (C, D): b200vq_linear_out; backward = b200vq_linear_in on W^T (grad_a),
    b200vq_weight_grad (grad_W as (C, D), grad_b = column sums of grad_y)."""

    @staticmethod
    def forward(ctx, a, weight, bias):
        lib = _lib.load()
        C, D = weight.shape
        aa = a.detach().to(torch.float32).contiguous()
        N = aa.numel() // D
        y = torch.empty(*a.shape[:-1], C, dtype=torch.float32, device=a.device)
        with _on_device(a.device):
            _lib.check(lib.b200vq_linear_out(aa.data_ptr(), N, D, weight.detach().contiguous().data_ptr(), 0,
                                             _ptr(bias.detach() if bias is not None else None), C, y.data_ptr(), C, _stream()),
                       "b200vq_linear_out")
        ctx.save_for_backward(aa, weight)
        ctx.has_bias = bias is not None
        ctx.a_shape = tuple(a.shape)
        return y

    @staticmethod
    def backward(ctx, gy):
        aa, weight = ctx.saved_tensors
        lib = _lib.load()
        C, D = weight.shape
        g, ldg, N = _rows2d(gy.to(torch.float32))
        ga = gw = gb = None
        with _on_device(gy.device):
            if (ctx.needs_input_grad[0] and ctx.needs_input_grad[1] and lib.b200vq_grad_out_supported(C, D)):
                # one pass over grad_output: grad_a, grad_W and grad_b together
                ga = torch.empty(ctx.a_shape, dtype=torch.float32, device=gy.device)
                gw = torch.empty(C, D, dtype=torch.float32, device=gy.device)
                gb = torch.empty(C, dtype=torch.float32, device=gy.device) if ctx.has_bias else None
                ws = _wg_workspace(lib, C, D, gy.device)
                _lib.check(lib.b200vq_grad_out(g.data_ptr(), ldg, aa.data_ptr(), weight.detach().contiguous().data_ptr(), N, C, D,
                                               None, ga.data_ptr(), gw.data_ptr(), _ptr(gb), ws.data_ptr(), ws.numel(), _stream()),
                           "b200vq_grad_out")
                return ga, gw, gb
            if ctx.needs_input_grad[0]:
                wt = weight.detach().t().contiguous()   # (D, C)
                ga = torch.empty(ctx.a_shape, dtype=torch.float32, device=gy.device)
                _lib.check(lib.b200vq_linear_in(g.data_ptr(), ldg, N, C, wt.data_ptr(), None, D, ga.data_ptr(), _stream()),
                           "b200vq_linear_in(grad_a)")
            if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
                gw = torch.empty(C, D, dtype=torch.float32, device=gy.device)
                gb = torch.empty(C, dtype=torch.float32, device=gy.device) if ctx.has_bias else None
                ws = _wg_workspace(lib, C, D, gy.device)
                _lib.check(lib.b200vq_weight_grad_ws(g.data_ptr(), ldg, aa.data_ptr(), N, C, D, gw.data_ptr(), 0, _ptr(gb),
                                                     ws.data_ptr(), ws.numel(), _stream()), "b200vq_weight_grad_ws(out_linear)")
        return ga, gw, gb


def linear_in(x, weight, bias):
    """in_linear through the streaming kernels when supported, else F.linear."""
    D, C = weight.shape
    if linears_supported(x, weight, C, D):
        return _LinearIn.apply(x, weight, bias)
    return torch.nn.functional.linear(x, weight, bias)


def linear_out(a, weight, bias):
    """out_linear through the streaming kernels when supported, else F.linear."""
    C, D = weight.shape
    if linears_supported(a, weight, C, D):
        return _LinearOut.apply(a, weight, bias)
    return torch.nn.functional.linear(a, weight, bias)


def codebook_lookup(indices: torch.Tensor, weight: torch.Tensor, l2norm: bool, shape=None) -> torch.Tensor:
    """get_codebook_entry (bottleneck.py:327-341) through b200vq_decode."""
    _require_cuda(indices, "indices")
    _require_cuda(weight, "embedding.weight")
    lib = _lib.load()
    idx = indices.reshape(-1).to(torch.int64).contiguous()
    w = weight.detach().contiguous()
    K, D = w.shape
    out = torch.empty(idx.numel(), D, dtype=torch.float32, device=w.device)
    with _on_device(w.device):
        _lib.check(lib.b200vq_decode(idx.data_ptr(), idx.numel(), w.data_ptr(), K, D, int(l2norm), out.data_ptr(),
                                     None, _stream()), "b200vq_decode")
    return out.reshape(shape) if shape is not None else out.reshape(*indices.shape, D)


def decode_project(indices: torch.Tensor, weight: torch.Tensor, l2norm: bool, w_out: torch.Tensor, b_out=None,
                   add=None) -> torch.Tensor:
    """Bottleneck.decode (bottleneck.py:166-168) in one pass through b200vq_decode_project:
    (..., n) indices -> (..., n, C) = normalise(weight[idx]) w_out^T + b_out (+ add, an (n, C) table broadcast
    over the leading dims: the decoder's latent positional embedding, models/larp_tokenizer.py:463-464).
    Inference only (no autograd graph), like codebook_lookup."""
    _require_cuda(indices, "indices")
    _require_cuda(weight, "embedding.weight")
    lib = _lib.load()
    idx = indices.reshape(-1).to(torch.int64).contiguous()
    w = weight.detach().contiguous()
    wo = w_out.detach().contiguous()
    K, D = w.shape
    C = wo.shape[0]
    add_rows = 0
    if add is not None:
        add = add.detach().to(torch.float32).reshape(-1, C).contiguous()
        add_rows = add.shape[0]
        if indices.dim() < 1 or indices.shape[-1] != add_rows:
            raise ValueError(f"add has {add_rows} rows but indices end in {tuple(indices.shape)}")
    out = torch.empty(*indices.shape, C, dtype=torch.float32, device=w.device)
    with _on_device(w.device):
        _lib.check(lib.b200vq_decode_project(idx.data_ptr(), idx.numel(), w.data_ptr(), K, D, int(l2norm), wo.data_ptr(),
                                             _ptr(b_out.detach() if b_out is not None else None), C, _ptr(add), add_rows,
                                             out.data_ptr(), C, None, _stream()), "b200vq_decode_project")
    return out


def usage_stats(hist_first: torch.Tensor, hist_batch: torch.Tensor) -> torch.Tensor:
    """(4,) fp32 device tensor: index_usage, index_usage_batch, kl_uni, perplexity
    (trainers/larp_tokenizer_trainer.py:335-352), no host sync."""
    _require_cuda(hist_first, "hist_first")
    lib = _lib.load()
    out = torch.empty(4, dtype=torch.float32, device=hist_first.device)
    with _on_device(hist_first.device):
        _lib.check(lib.b200vq_usage_stats(hist_first.data_ptr(), hist_batch.data_ptr(), hist_first.numel(),
                                          out.data_ptr(), _stream()), "b200vq_usage_stats")
    return out


# device index -> (flat gradient buffer, number of gradient floats, (2, K) int32 histograms, factor the histogram
# tail was pre-multiplied by, data pointers of the gradient pieces) of the last one-call backward, for
# sharded.token_sharded_sync
LAST_FLAT: dict = {}


def _dist_world() -> int:
    import torch.distributed as dist
    return dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1


def _dist_backend() -> str:
    import torch.distributed as dist
    return str(dist.get_backend()) if (dist.is_available() and dist.is_initialized()) else ""


class _BottleneckFunction(torch.autograd.Function):
    """Whole Bottleneck step (in_linear -> quantizer -> out_linear + usage statistics) in ONE host
    call per direction: b200vq_bottleneck_forward / b200vq_bottleneck_backward.
    Outputs: output, projected_z, znorm, emb, ste, idx, losses(3), hist, hist_first, usage(4)."""

    @staticmethod
    def forward(ctx, x, w_in, b_in, weight, w_out, b_out, cfg):
        mode, l2norm, t_inv, beta, w_cb, n_first, seed, offset, philox_dev, act_dtype, add = cfg
        ctx.set_materialize_grads(False)   # unused outputs arrive as None, not as zero tensors
        lib = _lib.load()
        dev = x.device
        D, C = w_in.shape
        K = weight.shape[0]
        xx, ldx, N = _rows2d(x.detach())
        lead = x.shape[:-1]
        f32 = dict(dtype=torch.float32, device=dev)
        with _on_device(dev):
            out = torch.empty(*lead, C, dtype=act_dtype, device=dev)
            # one allocation for the fp32 (N, D) tensors: [projected_z |] znorm | ste
            if act_dtype is torch.float32:
                nd = torch.empty(3, *lead, D, **f32)
                z, znorm, ste = nd[0], nd[1], nd[2]
            else:   # projected_z in the autocast dtype (models/bottleneck.py:140-142 under AMP)
                nd = torch.empty(2, *lead, D, **f32)
                z, znorm, ste = torch.empty(*lead, D, dtype=act_dtype, device=dev), nd[0], nd[1]
            add_rows = 0
            if add is not None:
                add = add.detach().to(torch.float32).reshape(-1, C).contiguous()
                add_rows = add.shape[0]
            emb = torch.empty(K, D, **f32)
            small = torch.empty(K + N + 12, **f32)          # wnorm | zlen | losses(3) usage(4) input norms(2)
            wnorm, zlen, losses, usage = small[:K], small[K:K + N], small[K + N:K + N + 3], small[K + N + 3:K + N + 7]
            norms = small[K + N + 8:K + N + 10]
            idx = torch.empty(N, dtype=torch.int64, device=dev)
            hists = torch.empty(2, K, dtype=torch.int32, device=dev)
            ws = _workspace(dev, lib.b200vq_bottleneck_workspace_bytes(N, C, D, K))
            use_tc = 1 if SCORING == "tc" else 0
            rc = lib.b200vq_bottleneck_forward(
                xx.data_ptr(), ldx, N, C, w_in.data_ptr(), _ptr(b_in), weight.data_ptr(), K, D, w_out.data_ptr(), _ptr(b_out),
                mode, int(l2norm), t_inv, seed, offset, n_first, beta, w_cb, use_tc, z.data_ptr(), emb.data_ptr(),
                wnorm.data_ptr(), znorm.data_ptr(), ste.data_ptr(), idx.data_ptr(), zlen.data_ptr(), hists[0].data_ptr(),
                hists[1].data_ptr(), losses.data_ptr(), usage.data_ptr(), norms.data_ptr(), out.data_ptr(), C, ws.data_ptr(), ws.numel(),
                _ptr(philox_dev), _DTYPES[act_dtype], _ptr(add), add_rows, _stream())
            _lib.check(rc, "b200vq_bottleneck_forward")
        ctx.save_for_backward(x, w_in, w_out, emb, wnorm, znorm, zlen, idx, ste)
        ctx.cfg = (bool(l2norm), beta, w_cb, N, C, D, K, b_in is not None, b_out is not None, act_dtype)
        ctx.hists = hists
        ctx.mark_non_differentiable(idx, hists, usage, norms)
        # the three losses leave as separate 0-dim outputs: indexing a (3,) output in the caller would put a
        # SelectBackward node (zeros + copy kernels) per used loss into every backward
        return out, z, znorm, emb, ste, idx, losses[0], losses[1], losses[2], hists, usage, norms

    @staticmethod
    def backward(ctx, g_out, g_z, g_znorm, g_emb, g_ste, _gi, g_lq, g_lc, g_lcb, _gh, _gu, _gn):
        x, w_in, w_out, emb, wnorm, znorm, zlen, idx, ste = ctx.saved_tensors
        l2norm, beta, w_cb, N, C, D, K, has_b_in, has_b_out, act_dtype = ctx.cfg
        lib = _lib.load()
        dev = x.device

        def _f32c(t):
            if t is None or (t.dtype is torch.float32 and t.is_contiguous()):
                return t
            return t.to(torch.float32).contiguous()

        g_z, g_znorm, g_emb, g_ste = _f32c(g_z), _f32c(g_znorm), _f32c(g_emb), _f32c(g_ste)
        g_lq, g_lc, g_lcb = _f32c(g_lq), _f32c(g_lc), _f32c(g_lcb)
        go, ldg = None, C
        if g_out is not None:
            # grad of 'output' arrives in the output's dtype (bf16 / fp16 under autocast) and is read as such
            go, ldg, _ = _rows2d(g_out if g_out.dtype is act_dtype else g_out.to(act_dtype))
        need = ctx.needs_input_grad
        f32 = dict(dtype=torch.float32, device=dev)
        with _on_device(dev):
            xx, ldx, _ = _rows2d(x.detach())
            gx = torch.empty(x.shape, **f32) if need[0] else None
            # every parameter gradient is a view of ONE flat buffer [codebook | W_in | b_in | W_out | b_out]
            # (16-byte aligned pieces), so a token-sharded job all-reduces them with one collective and no
            # packing copies (sharded.token_sharded_sync)
            sizes = [K * D, D * C, D, C * D, C]
            offs, tot = [], 0
            for n in sizes:
                offs.append(tot)
                tot += (n + 3) // 4 * 4
            sync = _dist_world() > 1
            hists = ctx.hists if sync else None
            flat = torch.empty(tot + (hists.numel() if sync else 0), **f32)
            if sync:   # usage histograms ride in the same collective (counts are exact in fp32 below 2^24)
                # pre-multiplied by the world size when it is a power of two: one all-reduce(AVG) then yields
                # averaged gradients AND summed histograms, no separate divide kernel
                world = _dist_world()
                scale = world if (world & (world - 1)) == 0 and _dist_backend() == "nccl" else 1
                torch.mul(hists.view(-1), scale, out=flat[tot:])
            gw = flat[offs[0]:offs[0] + sizes[0]].view(K, D) if need[3] else None
            gw_in = flat[offs[1]:offs[1] + sizes[1]].view(D, C) if need[1] else None
            gb_in = flat[offs[2]:offs[2] + sizes[2]] if (has_b_in and need[2]) else None
            gw_out = flat[offs[3]:offs[3] + sizes[3]].view(C, D) if (need[4] or (has_b_out and need[5])) else None
            gb_out = flat[offs[4]:offs[4] + sizes[4]] if (has_b_out and need[5]) else None
            if sync:
                ptrs = frozenset(t.data_ptr() for t in (gw, gw_in, gb_in, gw_out, gb_out) if t is not None)
                LAST_FLAT[dev.index] = (flat, tot, hists, scale, ptrs)
            if tot != sum(sizes) or any(t is None for t in (gw, gw_in, gb_in, gw_out, gb_out)):
                flat[:tot].zero_()   # padding / unused pieces must not carry garbage into the all-reduce
            g_losses = None
            if g_lq is not None or g_lc is not None or g_lcb is not None:
                g_losses = torch.empty(3, **f32)
                _lib.check(lib.b200vq_pack3(_ptr(g_lq), _ptr(g_lc), _ptr(g_lcb), g_losses.data_ptr(), _stream()), "b200vq_pack3")
            ws = _workspace(dev, lib.b200vq_bottleneck_workspace_bytes(N, C, D, K))
            # data-parallel jobs: an event recorded when the parameter gradients are enqueued, before the grad_x pass
            # (sharded.token_sharded_sync starts its all-reduce behind it on another stream)
            ready = GRADS_READY.get(dev.index) if sync else None
            rc = lib.b200vq_bottleneck_backward(
                _ptr(go), _DTYPES[act_dtype], ldg, _ptr(g_losses), _ptr(g_ste), _ptr(g_znorm), _ptr(g_emb), _ptr(g_z), xx.data_ptr(), ldx, N, C,
                w_in.data_ptr(), w_out.data_ptr(), emb.data_ptr(), wnorm.data_ptr(), znorm.data_ptr(), zlen.data_ptr(),
                idx.data_ptr(), ste.data_ptr(), K, D, int(l2norm), beta, w_cb, _ptr(gx), C, _ptr(gw_in), _ptr(gb_in), _ptr(gw),
                _ptr(gw_out), _ptr(gb_out), ws.data_ptr(), ws.numel(), ready.cuda_event if ready is not None else None, _stream())
            _lib.check(rc, "b200vq_bottleneck_backward")
        return gx, gw_in, gb_in, gw, gw_out, gb_out, None


def autocast_act_dtype() -> torch.dtype:
    """dtype the Bottleneck's linears hand back: the CUDA autocast dtype inside an autocast region, else fp32"""
    if torch.is_autocast_enabled():
        return torch.get_autocast_dtype("cuda") if hasattr(torch, "get_autocast_dtype") else torch.get_autocast_gpu_dtype()
    return torch.float32


def bottleneck_supported(x, w_in, weight, w_out) -> bool:
    """The one-call path needs fp32 x and parameters (under autocast the activations it hands back are bf16 / fp16, like
    the reference's nn.Linears), contiguous parameters and shapes the streaming linear kernels take."""
    D, C = w_in.shape
    act = autocast_act_dtype()
    lib = _lib.load()
    return (LINEARS == "fused" and x.is_cuda and x.dim() == 3 and x.dtype == torch.float32
            and w_in.dtype == weight.dtype == w_out.dtype == torch.float32
            and w_in.is_contiguous() and weight.is_contiguous() and w_out.is_contiguous()
            and tuple(w_out.shape) == (w_out.shape[0], D) and weight.shape[1] == D
            and act in _DTYPES and bool(lib.b200vq_linear_supported(C, D))
            and (act is torch.float32 or bool(lib.b200vq_grad_out_supported(C, D)))
            and bool(lib.b200vq_linear_supported(w_out.shape[0], D)) and w_out.shape[0] == C)


def bottleneck_step(x, w_in, b_in, weight, w_out, b_out, *, mode, l2norm, t_inv, beta, w_cb, n_first, seed, offset,
                    philox_dev=None, add=None):
    """-> (output, projected_z, znorm, emb, ste, idx(N,), loss_q, loss_commit, loss_codebook (0-dim each),
    hists(2,K) [batch, first], usage(4), input norms(2)).  philox_dev: optional int64[2] DEVICE tensor {seed, offset}
    read (and advanced) on the GPU instead of `seed` / `offset` -- what makes a captured CUDA graph replayable.
    Inside torch.autocast, output / projected_z come back in the autocast dtype.  add: optional (n, C) table added to
    `output` in the out_linear store (the decoder's latent positional embedding, models/larp_tokenizer.py:463-464); it is
    treated as a constant (a buffer in every shipped config: learned_decoder_latent_pe = False)."""
    cfg = (int(mode), bool(l2norm), float(t_inv), float(beta), float(w_cb), int(n_first), int(seed), int(offset), philox_dev,
           autocast_act_dtype(), add)
    return _BottleneckFunction.apply(x, w_in, b_in, weight, w_out, b_out, cfg)

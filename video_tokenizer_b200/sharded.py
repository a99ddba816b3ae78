"""Multi-GPU modes of the quantizer.  One process per GPU, torch.distributed
(NCCL over NVLink on the box; gloo in the CPU tests of the host logic).

Token-sharded (default, matches the reference's DDP: trainers/base_trainer.py:316,388):
    tokens are independent given the codebook, so the forward needs no collective.  After
    the backward, ONE all-reduce carries [codebook grad | linear grads | usage histograms].

Codebook-sharded (large-K sweeps): rank r owns codes [r*K/G, (r+1)*K/G).  Each rank scores
    its slice (b200vq_scores_partial), ONE all-reduce(MAX) of the 64-bit keys picks the
    global winner (highest score, lowest index), ONE all-reduce(SUM) of owner-only rows
    delivers the selected code vectors; the backward scatter-adds only owned rows, so the
    codebook gradient needs no collective.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import os

import torch
import torch.distributed as dist

SIGN = -(2 ** 63)


# ---- key algebra (pure tensor ops: also exercised by the CPU gloo tests) -------------------
def keys_to_signed(keys: torch.Tensor) -> torch.Tensor:
    """uint64 keys are carried in int64 tensors; flipping the top bit makes a SIGNED max
    equal to the unsigned max, which every backend (NCCL, gloo) supports for int64."""
    return keys ^ SIGN


def keys_from_signed(keys: torch.Tensor) -> torch.Tensor:
    return keys ^ SIGN


def key_indices(keys: torch.Tensor) -> torch.Tensor:
    """low 32 bits hold ~code"""
    return (~keys) & 0xFFFFFFFF


def allreduce_max_keys(keys: torch.Tensor, group=None) -> torch.Tensor:
    """In-place cross-rank max-loc on packed keys: one collective of 8*N bytes."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        s = keys_to_signed(keys)
        dist.all_reduce(s, op=dist.ReduceOp.MAX, group=group)
        keys.copy_(keys_from_signed(s))
    return keys


def shard_bounds(K: int, world: int, rank: int, align: int = 8) -> Tuple[int, int]:
    """[begin, end) of the code slice rank owns; slice starts are multiples of ``align``
    (sampling noise is generated for groups of 8 consecutive codes)."""
    per = -(-K // world)
    per = -(-per // align) * align
    b = min(K, rank * per)
    return b, min(K, b + per)


# ---- token-sharded gradient / statistics exchange -------------------------------------------
def fused_allreduce(tensors: Sequence[torch.Tensor], average: Sequence[bool], group=None) -> None:
    """One all-reduce over a single contiguous fp32 buffer holding every tensor; tensors
    flagged ``average`` are divided by the world size (DDP semantics), the others (integer
    histograms, exact in fp32 below 2^24 counts) are summed.  Results are written back."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    world = dist.get_world_size(group)
    flat = torch.cat([t.reshape(-1).to(torch.float32) for t in tensors])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    off = 0
    for t, avg in zip(tensors, average):
        n = t.numel()
        piece = flat[off:off + n].view_as(t)
        if avg:
            piece = piece / world
        t.copy_(piece.to(t.dtype))
        off += n


def _same_storage(a: torch.Tensor, b: torch.Tensor) -> bool:
    return a.untyped_storage().data_ptr() == b.untyped_storage().data_ptr()


def _flat_of(grads: Sequence[torch.Tensor]):
    """(flat, n_grad_floats, hists or None, factor the histogram tail was pre-multiplied by) if every gradient is a
    piece of one flat fp32 buffer, else (None, 0, None, 1)."""
    g0 = grads[0]
    if g0.is_cuda:
        from . import functional as VF
        packed = VF.LAST_FLAT.get(g0.device.index)
        if packed is not None and all(g.data_ptr() in packed[4] for g in grads):
            return packed[:4]
    base = g0._base
    if (base is not None and base.dim() == 1 and base.is_contiguous() and base.dtype == torch.float32
            and all(g._base is base for g in grads)):
        return base, base.numel(), None, 1
    return None, 0, None, 1


_SYNC_STREAMS: dict = {}


def enable_grad_overlap(device=None) -> None:
    """Let `token_sharded_sync` start its all-reduce while the backward's grad_x pass is still running: the one-call
    backward then records an event once the parameter gradients are enqueued (b200vq_bottleneck_backward,
    `grads_ready_event`), and the sync waits for that event on a stream of its own instead of for the whole backward.
    Call once per process after torch.cuda.set_device; `disable_grad_overlap()` goes back to the in-order sync."""
    from . import functional as VF
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    ev = torch.cuda.Event()
    with torch.cuda.device(dev):
        ev.record()          # creates the underlying cudaEvent_t (torch makes it lazily)
    VF.GRADS_READY[dev.index] = ev
    _SYNC_STREAMS.setdefault(dev.index, torch.cuda.Stream(device=dev))


def disable_grad_overlap(device=None) -> None:
    from . import functional as VF
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    VF.GRADS_READY.pop(dev.index, None)


def _ready_event(flat: torch.Tensor):
    if not flat.is_cuda:
        return None
    from . import functional as VF
    return VF.GRADS_READY.get(flat.device.index)


def _sync_stream(device) -> "torch.cuda.Stream":
    return _SYNC_STREAMS[device.index]


def token_sharded_sync(bottleneck, group=None) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
    """After loss.backward() on every rank: average the bottleneck's parameter grads and sum
    the fused usage histograms across ranks in ONE collective.  Returns the global
    (hist_first, hist_batch) or None if there is no regularizer state."""
    params = [p for p in bottleneck.parameters() if p.grad is not None]
    tensors: List[torch.Tensor] = [p.grad for p in params]
    flags = [True] * len(tensors)
    usage = getattr(getattr(bottleneck, "regularizer", None), "last_usage", None)
    # fast path: the one-call backward hands out every parameter gradient as a piece of one flat buffer
    # (autograd keeps the storage, so p.grad aliases it): ONE collective over [grads | histograms as fp32]
    if tensors and dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        world = dist.get_world_size(group)
        flat, tot, hists, scale = _flat_of(tensors)
        if flat is not None:
            tail_ok = hists is not None and usage is not None and _same_storage(usage[0], hists)
            if tail_ok and scale == world:
                # histograms were pre-multiplied by the world size: AVG gives mean gradients and summed counts
                ready = _ready_event(flat)
                if ready is not None:
                    # the one-call backward recorded `ready` when the parameter gradients were enqueued and went on to
                    # write grad_x: the collective runs beside that pass on its own stream
                    cur, side = torch.cuda.current_stream(flat.device), _sync_stream(flat.device)
                    side.wait_event(ready)
                    with torch.cuda.stream(side):
                        dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=group)
                        hists.view(-1).copy_(flat[tot:])
                    cur.wait_stream(side)
                    return usage
                dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=group)
                hists.view(-1).copy_(flat[tot:])
                return usage
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
            if tail_ok:
                if scale != 1:
                    flat[tot:].div_(scale)
                hists.view(-1).copy_(flat[tot:])
                flat[:tot].div_(world)
                return usage
            flat.div_(world)
            if usage is not None:
                hf, hb = usage
                if _same_storage(hf, hb) and hf._base is not None and hf._base.is_contiguous():
                    dist.all_reduce(hf._base, op=dist.ReduceOp.SUM, group=group)   # (2, K) int32: one more small collective
                else:
                    fused_allreduce([hf, hb], [False, False], group)
            return usage
    if usage is not None:
        hf, hb = usage
        if int(hb.numel()) and dist.is_initialized():
            assert hb.dtype == torch.int32
        tensors += [hf, hb]
        flags += [False, False]
    fused_allreduce(tensors, flags, group)
    return usage


# ---- codebook-sharded forward ---------------------------------------------------------------
def codebook_sharded_select(score_slice: Callable[[int, int], torch.Tensor],
                            gather_owned: Callable[[torch.Tensor, int, int], torch.Tensor],
                            K: int, group=None):
    """Host logic of the codebook-sharded forward, with the two device steps injected:
        score_slice(k_begin, k_end)       -> keys (N,) int64 carrying uint64 keys for that slice
        gather_owned(keys, k_begin, k_end) -> (N, D) fp32 rows, zero where the code is not owned
    Returns (global keys, global indices (N,), selected code rows (N, D))."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    kb, ke = shard_bounds(K, world, rank)
    keys = score_slice(kb, ke)
    allreduce_max_keys(keys, group)
    rows = gather_owned(keys, kb, ke)
    if world > 1:
        dist.all_reduce(rows, op=dist.ReduceOp.SUM, group=group)
    return keys, key_indices(keys), rows


# ---- codebook-sharded quantizer on the device (C ABI + NCCL) ----------------------------------
class _CodebookShardedVQ(torch.autograd.Function):
    """Quantizer step with the codebook rows [k_begin, k_begin + K_local) on this rank and the tokens
    replicated: local scores -> all-reduce(MAX) of the packed keys -> owner-only
    rows all-reduce(SUM) -> finish.  The backward scatter-adds owned rows only, so the codebook
    gradient needs no collective; grad_z is identical on every rank."""

    @staticmethod
    def forward(ctx, z, weight_local, cfg):
        from . import _lib, functional as VF
        K, k_begin, mode, l2norm, t_inv, beta, w_cb, n_first, seed, offset, group = cfg
        ctx.set_materialize_grads(False)
        lib = _lib.load()
        dev = z.device
        D = z.shape[-1]
        K_local = weight_local.shape[0]
        zz, stride = VF.rows_view(z.detach())
        N = zz.shape[:-1].numel()
        st = VF._stream
        f32 = dict(dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            emb_l, e2_l, wnorm_l = VF.prepare_codebook(weight_local, l2norm)
            # Equal shards: gather the normalised slices into a full (K, D) table WHILE the local slice is scored (the
            # table does not depend on the scores); the winning rows are then read locally and the (N, D) all-reduce of
            # the owners' rows is not needed.  Unequal shards keep the owner-rows exchange.
            world = dist.get_world_size(group) if dist.is_initialized() else 1
            gather_work, emb_full = None, None
            if world > 1 and K % world == 0 and K_local == K // world and k_begin == dist.get_rank(group) * K_local \
                    and os.environ.get("B200VQ_SHARDED_ROWS", "gather") == "gather":
                emb_full = torch.empty(K, D, **f32)
                gather_work = dist.all_gather_into_tensor(emb_full, emb_l, group=group, async_op=True)
            keys = torch.zeros(N, dtype=torch.int64, device=dev)
            use_tc = VF.SCORING == "tc" and lib.b200vq_tc_supported(D, int(l2norm), mode) and k_begin % 256 == 0
            if use_tc:
                image = VF.prepare_codebook_tc(emb_l)
                scratch = torch.zeros(lib.b200vq_tc_scratch_bytes(N) // 4 + 4, **f32)
                _lib.check(lib.b200vq_scores_tc(zz.data_ptr(), VF._DTYPES[zz.dtype], stride, N, D, image.data_ptr(),
                                                emb_l.data_ptr(), e2_l.data_ptr(), K_local, k_begin, mode, int(l2norm),
                                                t_inv, seed, offset, 0, keys.data_ptr(), scratch.data_ptr(), None, st()),
                           "b200vq_scores_tc")
            else:
                _lib.check(lib.b200vq_scores_partial(zz.data_ptr(), VF._DTYPES[zz.dtype], stride, N, D, emb_l.data_ptr(),
                                                     e2_l.data_ptr(), K_local, k_begin, mode, int(l2norm), t_inv, seed,
                                                     offset, 0, keys.data_ptr(), st()), "b200vq_scores_partial")
            allreduce_max_keys(keys, group)                                   # the cross-GPU max-loc
            rows = torch.empty(N, D, **f32)
            if gather_work is not None:
                gather_work.wait()                                            # the current stream waits for the table
                _lib.check(lib.b200vq_gather_owned(keys.data_ptr(), N, emb_full.data_ptr(), K, 0, D, rows.data_ptr(), st()),
                           "b200vq_gather_owned")
            else:
                _lib.check(lib.b200vq_gather_owned(keys.data_ptr(), N, emb_l.data_ptr(), K_local, k_begin, D, rows.data_ptr(),
                                                   st()), "b200vq_gather_owned")
                if world > 1:
                    dist.all_reduce(rows, op=dist.ReduceOp.SUM, group=group)  # only the owner wrote non-zero
            znorm = torch.empty(*z.shape, **f32)
            ste = torch.empty(*z.shape, **f32)
            idx = torch.empty(N, dtype=torch.int64, device=dev)
            zlen = torch.empty(N, **f32)
            hists = torch.zeros(2, K, dtype=torch.int32, device=dev)
            sqerr = torch.zeros(2, dtype=torch.float64, device=dev)
            losses = torch.empty(3, **f32)
            _lib.check(lib.b200vq_finish(zz.data_ptr(), VF._DTYPES[zz.dtype], stride, N, D, None, K, keys.data_ptr(),
                                         rows.data_ptr(), int(l2norm), n_first, znorm.data_ptr(), ste.data_ptr(),
                                         idx.data_ptr(), zlen.data_ptr(), hists[0].data_ptr(), hists[1].data_ptr(),
                                         sqerr.data_ptr(), 1, N, beta, w_cb, losses.data_ptr(), st()), "b200vq_finish")
        ctx.save_for_backward(znorm, zlen, idx, emb_l, wnorm_l, rows)
        ctx.cfg = (bool(l2norm), beta, w_cb, N, D, K_local, k_begin, z.dtype, tuple(z.shape))
        ctx.mark_non_differentiable(idx, hists)
        return znorm, ste, idx, losses, hists

    @staticmethod
    def backward(ctx, g_znorm, g_ste, _gi, g_losses, _gh):
        from . import _lib, functional as VF
        znorm, zlen, idx, emb_l, wnorm_l, rows = ctx.saved_tensors
        l2norm, beta, w_cb, N, D, K_local, k_begin, z_dtype, z_shape = ctx.cfg
        lib = _lib.load()
        dev = znorm.device

        def _f32c(t):
            return None if t is None else t.to(torch.float32).contiguous()

        g_znorm, g_ste, g_losses = _f32c(g_znorm), _f32c(g_ste), _f32c(g_losses)
        with torch.cuda.device(dev):
            gz = torch.empty(z_shape, dtype=z_dtype, device=dev)
            grad_emb = torch.empty(K_local, D, dtype=torch.float32, device=dev)
            _lib.check(lib.b200vq_backward(VF._ptr(g_ste), VF._ptr(g_znorm), VF._ptr(g_losses), znorm.data_ptr(),
                                           zlen.data_ptr(), idx.data_ptr(), emb_l.data_ptr(), rows.data_ptr(), k_begin, N, D,
                                           K_local, int(l2norm), N, beta, w_cb, gz.data_ptr(), VF._DTYPES[z_dtype],
                                           grad_emb.data_ptr(), VF._stream()), "b200vq_backward")
            gw = torch.empty(K_local, D, dtype=torch.float32, device=dev)
            _lib.check(lib.b200vq_codebook_backward(grad_emb.data_ptr(), None, emb_l.data_ptr(), wnorm_l.data_ptr(), K_local, D,
                                                    int(l2norm), gw.data_ptr(), VF._stream()), "b200vq_codebook_backward")
        return gz, gw, None


def codebook_sharded_quantize(z, weight_local, K, k_begin, *, mode, l2norm, t_inv=1.0, beta=0.25, w_cb=1.0, n_first=None,
                              seed=0, offset=0, group=None):
    """Codebook-sharded quantizer step.  ``weight_local`` holds rows [k_begin, k_begin + K_local) of the
    (K, D) codebook (``shard_bounds(K, world, rank, align=256)`` keeps the tensor-core path usable); z is
    the same on every rank; in stochastic mode every rank must pass the same (seed, offset).
    Returns (znorm, ste, idx (N,), losses (3), hists (2, K) int32 [batch, first instance])."""
    if n_first is None:
        n_first = z.shape[1] if z.dim() == 3 else z.shape[0]
    cfg = (int(K), int(k_begin), int(mode), bool(l2norm), float(t_inv), float(beta), float(w_cb), int(n_first), int(seed),
           int(offset), group)
    return _CodebookShardedVQ.apply(z, weight_local, cfg)

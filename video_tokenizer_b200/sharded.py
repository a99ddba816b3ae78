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


def token_sharded_sync(bottleneck, group=None) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
    """After loss.backward() on every rank: average the bottleneck's parameter grads and sum
    the fused usage histograms across ranks in ONE collective.  Returns the global
    (hist_first, hist_batch) or None if there is no regularizer state."""
    params = [p for p in bottleneck.parameters() if p.grad is not None]
    tensors: List[torch.Tensor] = [p.grad for p in params]
    flags = [True] * len(tensors)
    usage = getattr(getattr(bottleneck, "regularizer", None), "last_usage", None)
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

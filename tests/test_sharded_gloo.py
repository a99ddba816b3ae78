"""world_size-2 gloo tests (CPU) of the multi-GPU host logic in video_tokenizer_b200.sharded:
key algebra, shard bounds, the codebook-sharded select protocol and the token-sharded fused
all-reduce.  The device steps are injected as plain torch CPU stand-ins built from the oracle,
so what is tested here is the exchange protocol, not a compute path."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from video_tokenizer_b200 import sharded


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _orderable(v: torch.Tensor) -> torch.Tensor:
    """float32 -> monotone uint32 carried in int64 (device code: vq_common.cuh orderable())"""
    v = v + 0.0
    b = v.contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    neg = (b & 0x80000000) != 0
    return torch.where(neg, (~b) & 0xFFFFFFFF, b | 0x80000000)


def _make_keys(score: torch.Tensor, code: torch.Tensor) -> torch.Tensor:
    """(orderable(score) << 32) | ~code as the uint64 bit pattern inside an int64 tensor"""
    hi = _orderable(score)
    lo = (~code.to(torch.int64)) & 0xFFFFFFFF
    hi_signed = torch.where(hi >= 2 ** 31, hi - 2 ** 32, hi)   # top bit -> sign
    return (hi_signed << 32) | lo


def _problem(N=257, D=16, K=1000):
    g = torch.Generator().manual_seed(7)
    z = torch.nn.functional.normalize(torch.randn(N, D, generator=g), dim=-1)
    emb = torch.nn.functional.normalize(torch.randn(K, D, generator=g), dim=-1)
    emb[5] = emb[900]      # exact tie across the two shards -> lowest index must win
    z[0] = emb[900]
    return z, emb


def _worker(rank, world, port, fn, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ret[rank] = fn(rank, world)
    finally:
        dist.destroy_process_group()


def _run(fn, world=2):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, fn, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0, f"rank exited with {p.exitcode}"
    return dict(ret)


# ---- pure functions (no process group) ---------------------------------------------------------
def test_key_roundtrip_and_order():
    s = torch.tensor([-1.0, -0.0, 0.0, 0.5, 0.5, 1.0])
    c = torch.tensor([3, 9, 2, 7, 1, 0])
    k = _make_keys(s, c)
    assert torch.equal(sharded.key_indices(k), c)
    sk = sharded.keys_to_signed(k)
    assert torch.equal(sharded.keys_from_signed(sk), k)
    # signed order of the flipped keys == (score, then lowest code)
    assert sk[5] > sk[4] > sk[3] > sk[2] and sk[1] == _sk(s[2:3], c[1:2]) and sk[2] > sk[1] > sk[0]


def _sk(s, c):
    return sharded.keys_to_signed(_make_keys(s, c))[0]


@pytest.mark.parametrize("K,world", [(8192, 2), (1000, 2), (1000, 8), (7, 4), (262144, 8)])
def test_shard_bounds_cover_and_align(K, world):
    prev = 0
    for r in range(world):
        b, e = sharded.shard_bounds(K, world, r)
        assert b == prev and b <= e <= K and (b % 8 == 0 or b == K)
        prev = e
    assert prev == K


# ---- world_size 2 ------------------------------------------------------------------------------
def _select(rank, world):
    z, emb = _problem()
    K = emb.shape[0]

    def score_slice(kb, ke):
        s = z @ emb[kb:ke].t()
        best, arg = s.max(dim=1)                       # torch.max returns the first maximal index
        first = (s == best[:, None]).float().argmax(dim=1)
        return _make_keys(best, first + kb)

    def gather_owned(keys, kb, ke):
        idx = sharded.key_indices(keys)
        own = (idx >= kb) & (idx < ke)
        rows = torch.zeros(z.shape[0], emb.shape[1])
        rows[own] = emb[idx[own]]
        return rows

    keys, idx, rows = sharded.codebook_sharded_select(score_slice, gather_owned, K)
    return idx.clone(), rows.clone()


def test_codebook_sharded_select_equals_unsharded():
    out = _run(_select)
    z, emb = _problem()
    s = z @ emb.t()
    ref = (s == s.max(dim=1, keepdim=True).values).float().argmax(dim=1)
    assert int(ref[0]) == 5          # the duplicated code: lowest global index wins across shards
    for r in (0, 1):
        idx, rows = out[r]
        assert torch.equal(idx, ref)
        assert torch.equal(rows, emb[ref])


def _fused(rank, world):
    g = torch.Generator().manual_seed(100 + rank)
    grads = [torch.randn(8, 4, generator=g), torch.randn(5, generator=g)]
    hist = torch.randint(0, 50, (16,), generator=g, dtype=torch.int32)
    mine = [t.clone() for t in grads] + [hist.clone()]
    sharded.fused_allreduce(grads + [hist], [True, True, False])
    return mine, [t.clone() for t in grads] + [hist.clone()]


def test_fused_allreduce_averages_grads_and_sums_histograms():
    out = _run(_fused)
    (m0, r0), (m1, r1) = out[0], out[1]
    for a, b in zip(r0, r1):
        assert torch.equal(a, b)
    assert torch.allclose(r0[0], (m0[0] + m1[0]) / 2) and torch.allclose(r0[1], (m0[1] + m1[1]) / 2)
    assert r0[2].dtype == torch.int32 and torch.equal(r0[2], m0[2] + m1[2])


class _FakeReg(torch.nn.Module):
    def __init__(self, K, rank):
        super().__init__()
        self.embedding = torch.nn.Embedding(K, 4)
        g = torch.Generator().manual_seed(rank)
        self.last_usage = (torch.randint(0, 9, (K,), generator=g, dtype=torch.int32),
                           torch.randint(0, 99, (K,), generator=g, dtype=torch.int32))


class _FakeBottleneck(torch.nn.Module):
    def __init__(self, rank):
        super().__init__()
        self.in_linear = torch.nn.Linear(6, 4)
        self.regularizer = _FakeReg(32, rank)


def _sync(rank, world):
    torch.manual_seed(0)
    m = _FakeBottleneck(rank)
    g = torch.Generator().manual_seed(10 + rank)
    for p in m.parameters():
        p.grad = torch.randn(p.shape, generator=g)
    before = [p.grad.clone() for p in m.parameters()]
    hb = [t.clone() for t in m.regularizer.last_usage]
    usage = sharded.token_sharded_sync(m)
    return before, [p.grad.clone() for p in m.parameters()], hb, [u.clone() for u in usage]


def test_token_sharded_sync_one_collective_semantics():
    out = _run(_sync)
    b0, a0, h0, u0 = out[0]
    b1, a1, h1, u1 = out[1]
    for x0, x1, y0, y1 in zip(b0, b1, a0, a1):
        assert torch.allclose(y0, (x0 + x1) / 2) and torch.equal(y0, y1)
    for i in range(2):
        assert torch.equal(u0[i], h0[i] + h1[i]) and torch.equal(u0[i], u1[i])


def _sync_flat(rank, world):
    """parameter grads handed out as views of one flat buffer + (2, K) histograms: the fast path"""
    torch.manual_seed(0)
    m = _FakeBottleneck(rank)
    g = torch.Generator().manual_seed(20 + rank)
    ps = list(m.parameters())
    flat = torch.randn(sum(p.numel() for p in ps) + 3, generator=g)
    off = 0
    for p in ps:
        p.grad = flat[off:off + p.numel()].view(p.shape)
        off += p.numel()
    hists = torch.randint(0, 99, (2, 32), generator=g, dtype=torch.int32)
    m.regularizer.last_usage = (hists[1], hists[0])
    before = [p.grad.clone() for p in ps]
    hb = hists.clone()
    usage = sharded.token_sharded_sync(m)
    return before, [p.grad.clone() for p in ps], hb, torch.stack([usage[1], usage[0]]).clone()


def test_token_sharded_sync_flat_buffer_fast_path():
    out = _run(_sync_flat)
    b0, a0, h0, u0 = out[0]
    b1, a1, h1, u1 = out[1]
    for x0, x1, y0, y1 in zip(b0, b1, a0, a1):
        assert torch.allclose(y0, (x0 + x1) / 2) and torch.equal(y0, y1)
    assert torch.equal(u0, h0 + h1) and torch.equal(u0, u1)

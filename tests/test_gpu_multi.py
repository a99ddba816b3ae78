"""Two-GPU tests (NCCL) of the sharded modes against the unsharded result on the same inputs.
Skipped unless the box has >= 2 CUDA devices (gpurun --gpus 2)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, name, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        ret[rank] = globals()[name](rank, world)
    finally:
        dist.destroy_process_group()


def _run(name, world=2):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} CUDA devices")
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, name, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0, f"rank exited with {p.exitcode}"
    return dict(ret)


def _codebook_sharded(rank, world):
    import video_tokenizer_b200 as V
    from video_tokenizer_b200 import _lib, functional as VF, sharded
    dev = torch.device("cuda", rank)
    out = {}
    for mode, K, D, N in [(_lib.MODE_ARGMIN_DIST, 4096, 16, 3000), (_lib.MODE_SAMPLE, 2048, 24, 1500),
                          (_lib.MODE_ARGMAX_COS, 1000, 8, 257)]:
        g = torch.Generator().manual_seed(5)
        z = torch.randn(3, N // 3, D, generator=g).to(dev).requires_grad_(True)
        W = (torch.rand(K, D, generator=g) - 0.5).to(dev).requires_grad_(True)
        G = torch.randn(3, N // 3, D, generator=g).to(dev)
        ref = VF.vq_quantize(z, W, mode=mode, l2norm=True, t_inv=1 / 0.03, seed=11, offset=8)
        torch.autograd.backward([ref[2], ref[4][0]], [G, torch.tensor(0.1, device=dev)])
        gz_ref, gw_ref = z.grad.clone(), W.grad.clone()
        z.grad = None
        kb, ke = sharded.shard_bounds(K, world, rank, align=256)
        Wl = W.detach()[kb:ke].clone().requires_grad_(True)
        znorm, ste, idx, losses, hists = sharded.codebook_sharded_quantize(z, Wl, K, kb, mode=mode, l2norm=True,
                                                                           t_inv=1 / 0.03, seed=11, offset=8)
        torch.autograd.backward([ste, losses[0]], [G, torch.tensor(0.1, device=dev)])
        torch.cuda.synchronize()
        out[mode] = dict(idx_equal=bool(torch.equal(idx, ref[3])), ste_equal=bool(torch.equal(ste, ref[2])),
                         loss_err=float((losses - ref[4]).abs().max()),
                         gz_err=float((z.grad - gz_ref).abs().max() / gz_ref.abs().max()),
                         gw_err=float((Wl.grad - gw_ref[kb:ke]).abs().max() / gw_ref.abs().max()),
                         hist_equal=bool(torch.equal(hists[0], ref[5])))
    return out


def test_codebook_sharded_equals_unsharded_two_gpus():
    out = _run("_codebook_sharded")
    for rank in (0, 1):
        for mode, r in out[rank].items():
            assert r["idx_equal"] and r["ste_equal"] and r["hist_equal"], (rank, mode, r)
            assert r["loss_err"] < 1e-6 and r["gz_err"] < 1e-5 and r["gw_err"] < 1e-5, (rank, mode, r)


def _token_sharded(rank, world):
    import video_tokenizer_b200 as V
    from video_tokenizer_b200 import sharded
    dev = torch.device("cuda", rank)
    torch.manual_seed(3)
    spec = {"name": "bottleneck", "args": {"bottleneck_dim": 16, "norm": "none", "regularizer": {"name": "vq", "args": {
        "codebook_size": 1024, "l2_normalized": True, "stochastic": False}}}}
    m = V.make(spec, args={"token_nums": 64, "input_dim": 128, "output_dim": 128}).to(dev)
    g = torch.Generator().manual_seed(50)
    xs = torch.randn(world, 4, 64, 128, generator=g)
    Gs = torch.randn(world, 4, 64, 128, generator=g)

    def run(x, G):
        for p in m.parameters():
            p.grad = None
        out = m(x.to(dev))
        torch.autograd.backward([out["output"], out["loss_q"]], [G.to(dev), torch.tensor(0.1, device=dev)])
        return [p.grad.clone() for p in m.parameters()], [h.clone() for h in m.regularizer.last_usage]

    # reference: both shards on this GPU, averaged by hand
    gs, hs = zip(*[run(xs[r], Gs[r]) for r in range(world)])
    ref_g = [sum(t) / world for t in zip(*gs)]
    ref_h = [sum(t) for t in zip(*hs)]
    run(xs[rank], Gs[rank])
    usage = sharded.token_sharded_sync(m)
    torch.cuda.synchronize()
    gerr = max(float((p.grad - r).abs().max() / (r.abs().max() + 1e-20)) for p, r in zip(m.parameters(), ref_g))
    res = dict(gerr=gerr, hist_equal=all(bool(torch.equal(a, b)) for a, b in zip(usage, ref_h)),
               fast_path=sharded._flat_of([p.grad for p in m.parameters()])[0] is not None)
    # the same sync started behind the backward's grads-ready event on its own stream (beside the grad_x pass),
    # eager and captured in a CUDA graph together with the step
    sharded.enable_grad_overlap(dev)
    run(xs[rank], Gs[rank])
    usage = sharded.token_sharded_sync(m)
    torch.cuda.synchronize()
    res["gerr_overlap"] = max(float((p.grad - r).abs().max() / (r.abs().max() + 1e-20)) for p, r in zip(m.parameters(), ref_g))
    res["hist_equal_overlap"] = all(bool(torch.equal(a, b)) for a, b in zip(usage, ref_h))
    step = V.GraphedBottleneckStep(m, xs[rank].to(dev), Gs[rank].to(dev), loss_q_grad=0.1, warmup=2,
                                   post_backward=lambda: sharded.token_sharded_sync(m))
    for _ in range(2):
        step.replay()
    torch.cuda.synchronize()
    res["gerr_graph"] = max(float((p.grad - r).abs().max() / (r.abs().max() + 1e-20)) for p, r in zip(m.parameters(), ref_g))
    res["hist_equal_graph"] = all(bool(torch.equal(a, b)) for a, b in zip(m.regularizer.last_usage, ref_h))
    sharded.disable_grad_overlap(dev)
    return res


def test_token_sharded_sync_two_gpus():
    out = _run("_token_sharded")
    for rank in (0, 1):
        assert out[rank]["gerr"] < 1e-5 and out[rank]["hist_equal"] and out[rank]["fast_path"], out[rank]
        assert out[rank]["gerr_overlap"] < 1e-5 and out[rank]["hist_equal_overlap"], out[rank]
        assert out[rank]["gerr_graph"] < 1e-5 and out[rank]["hist_equal_graph"], out[rank]

"""BASELINE config 5: codebook-size sweep 8K -> 256K codes, codebook-sharded max-loc over NVLink.

    python tools/ksweep.py                         # 1 GPU, unsharded
    torchrun --nproc-per-node G tools/ksweep.py    # G GPUs: codes sharded, tokens replicated

N = 65536 tokens, D = 16, deterministic (argmin) and stochastic; fwd+bwd of the quantizer step timed with
CUDA events (max over ranks).  --cpu adds the oracle (reference op sequence, all host cores) on a row sample.
Prints one JSON line per (K, mode) on rank 0.
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from video_tokenizer_b200 import _lib, functional as VF, sharded

ap = argparse.ArgumentParser()
ap.add_argument("--tokens", type=int, default=65536)
ap.add_argument("--dim", type=int, default=16)
ap.add_argument("--ks", default="8192,16384,32768,65536,131072,262144")
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--cpu", action="store_true", help="also time the CPU oracle on 1024 rows (rank 0)")
args = ap.parse_args()

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
lr = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
N, D = args.tokens, args.dim
g = torch.Generator().manual_seed(1234)
z = torch.randn(N // 1024, 1024, D, generator=g).to(dev).requires_grad_(True)
G = torch.randn(N // 1024, 1024, D, generator=g).to(dev)
gq = torch.tensor(0.1, device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

for K in [int(k) for k in args.ks.split(",")]:
    torch.manual_seed(0)
    W = torch.empty(K, D)
    torch.nn.init.kaiming_uniform_(W)
    kb, ke = sharded.shard_bounds(K, world, rank, align=256)
    Wl = W[kb:ke].to(dev).requires_grad_(True)
    for mode_name, mode in (("deterministic", _lib.MODE_ARGMIN_DIST), ("stochastic", _lib.MODE_SAMPLE)):
        def step():
            z.grad = None; Wl.grad = None
            if world > 1:
                zn, ste, idx, losses, hists = sharded.codebook_sharded_quantize(z, Wl, K, kb, mode=mode, l2norm=True,
                                                                               t_inv=1 / 0.03, seed=7, offset=16)
                loss = losses[0]
            else:
                zn, emb, ste, idx, losses, h, hf = VF.vq_quantize(z, Wl, mode=mode, l2norm=True, t_inv=1 / 0.03, seed=7,
                                                                 offset=16)
                loss = losses[0]
            torch.autograd.backward([ste, loss], [G, gq])
            return idx
        for _ in range(3):
            idx = step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0.record()
        for _ in range(args.reps):
            step()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.reps], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        line = {"config": "K sweep (BASELINE configs[4])", "K": K, "D": D, "tokens": N, "mode": mode_name, "n_gpus": world,
                "parallelism": "codebook-sharded" if world > 1 else "unsharded", "ms_fwd_bwd": ms,
                "tokens_per_s": N / (ms * 1e-3), "usage": int(torch.unique(idx).numel())}
        if args.cpu and rank == 0 and mode_name == "deterministic":
            from oracle import vq_oracle as O
            torch.set_num_threads(os.cpu_count() or 1)
            cfg = O.VQConfig(dim=D, codebook_size=K, l2_normalized=True)
            zc = z.detach().cpu()[:1].clone().requires_grad_(True)     # 1024 rows (row chunks do not change per-row results)
            Wc = W.clone().requires_grad_(True)
            t0 = time.perf_counter()
            out = O.vq_forward(zc, Wc, cfg)
            torch.autograd.backward([out["regularized_z"], out["loss_q"]], [G[:1].cpu(), torch.tensor(0.1)])
            dt = time.perf_counter() - t0
            same = bool(torch.equal(out["bottleneck_rep"].reshape(-1), idx[:1024].cpu()))
            line["cpu_reference"] = {"tokens_per_s": 1024 / dt, "cores": os.cpu_count(), "rows": 1024,
                                     "indices_equal_on_sample": same}
        if rank == 0:
            print(json.dumps(line), flush=True)
if world > 1:
    dist.destroy_process_group()

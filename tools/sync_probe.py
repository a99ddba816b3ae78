"""Tuning aid (torchrun, >= 2 GPUs): time token_sharded_sync and a bare all-reduce of the same size."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import video_tokenizer_b200 as V
from video_tokenizer_b200 import sharded

rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
torch.manual_seed(0)
spec = {"name": "bottleneck", "args": {"bottleneck_dim": 24, "norm": "none", "regularizer": {"name": "vq", "args": {
    "codebook_size": 8192, "l2_normalized": True, "stochastic": True, "stochastic_temperature": 0.03}}}}
m = V.make(spec, args={"token_nums": 1024, "input_dim": 768, "output_dim": 768}).to(dev)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
x = torch.randn(B, 1024, 768, device=dev, requires_grad=True)
G = torch.randn(B, 1024, 768, device=dev)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def step(sync):
    for p in m.parameters():
        p.grad = None
    out = m(x)
    torch.autograd.backward([out["output"], out["loss_q"]], [G, torch.tensor(0.1, device=dev)])
    if sync:
        sharded.token_sharded_sync(m)


def timeit(fn, reps=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


t0 = timeit(lambda: step(False)); t1 = timeit(lambda: step(True))
flat = torch.zeros(330000, device=dev)
t2 = timeit(lambda: dist.all_reduce(flat))
if rank == 0:
    print(f"step {t0:.1f} us, step+sync {t1:.1f} us, bare all_reduce(1.3 MB) {t2:.1f} us, "
          f"fast path {sharded._flat_of([p.grad for p in m.parameters()])[0] is not None}")
dist.destroy_process_group()

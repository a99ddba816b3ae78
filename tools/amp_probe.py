"""Kernel-level view of the bottleneck step under torch.autocast(bf16) against fp32 (torch.profiler, eager calls).
    python tools/amp_probe.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

import bench
import video_tokenizer_b200 as V


class A:
    mode = "stochastic"
    batch = 64


dev = torch.device("cuda")
V.load_library()
torch.manual_seed(0)
m = bench.make_module(V, A, dev)
x = torch.randn(64, 1024, 768, device=dev, requires_grad=True)
gq = torch.tensor(0.1, device=dev)


def step(dtype):
    for p in m.parameters():
        p.grad = None
    x.grad = None
    if dtype is None:
        out = m(x)
        G = G32
    else:
        with torch.autocast("cuda", dtype=dtype):
            out = m(x)
        G = G16
    torch.autograd.backward([out["output"], out["loss_q"]], [G, gq])


G32 = torch.randn(64, 1024, 768, device=dev)
G16 = G32.to(torch.bfloat16)
for name, dt in (("fp32", None), ("bf16", torch.bfloat16)):
    for _ in range(3):
        step(dt)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            step(dt)
        torch.cuda.synchronize()
    print(f"==== {name}")
    rows = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)
    tot = 0.0
    for e in rows[:22]:
        tot += e.device_time_total / 5
        print(f"{e.key[:70]:70s} n={e.count // 5:2d}  {e.device_time_total / 5:8.1f} us/step")
    print(f"sum of the listed kernels: {tot:.1f} us/step")

"""Streaming entropy regulariser (b200vq_entropy_forward / _backward) against the reference's N x K op sequence in
torch on the same device: fwd+bwd time and peak memory.  python tools/entropy_probe.py [out.json]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import video_tokenizer_b200 as V  # noqa: E402
from video_tokenizer_b200 import functional as VF  # noqa: E402


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters, torch.cuda.max_memory_allocated()


def main():
    out = []
    for N, K, D, T in [(4096, 8192, 24, 0.01), (65536, 8192, 24, 0.01), (65536, 8192, 16, 0.01), (16384, 65536, 16, 0.01)]:
        torch.manual_seed(0)
        z = torch.nn.functional.normalize(torch.randn(N, D, device="cuda"), dim=-1).requires_grad_(True)
        e = torch.nn.functional.normalize(torch.randn(K, D, device="cuda"), dim=-1).requires_grad_(True)

        def streamed():
            z.grad = e.grad = None
            VF.entropy_loss_streamed(z, e, T)[0].backward()

        def streamed_fwd():
            with torch.no_grad():
                VF.entropy_loss_streamed(z, e, T)

        def matrix():
            z.grad = e.grad = None
            d = torch.sum(z ** 2, dim=1, keepdim=True) + torch.sum(e ** 2, dim=1) - 2 * z @ e.t()
            V.entropy_loss(-d, temperature=T)[0].backward()

        base = torch.cuda.memory_allocated()
        ts, ms = timed(streamed)
        tf, _ = timed(streamed_fwd)
        tm, mm = timed(matrix, iters=3, warm=1)
        flops = 2.0 * N * K * D
        rec = {"N": N, "K": K, "D": D, "T": T, "streamed_fwd_bwd_ms": ts, "streamed_fwd_ms": tf, "torch_matrix_fwd_bwd_ms": tm,
               "streamed_peak_extra_MB": (ms - base) / 2**20, "torch_matrix_peak_extra_MB": (mm - base) / 2**20,
               "streamed_tflops_fp32": 5 * flops / ts / 1e9, "speedup": tm / ts}
        print(json.dumps(rec))
        out.append(rec)
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as f:
            for r in out:
                f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()

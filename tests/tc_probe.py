"""Bring-up probe for the tcgen05 path (run on the GPU box): tries both LBO/SBO descriptor
conventions and reports agreement with the exact path."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from video_tokenizer_b200 import _lib
from tests.test_gpu_tc import _scores

lib = _lib.load()
torch.manual_seed(0)
for swap in (0, 1):
    lib.b200vq_tc_debug_desc_swap(swap)
    for (D, K, N) in [(16, 256, 128), (24, 8192, 4096)]:
        z = torch.randn(N, D, device="cuda"); w = torch.randn(K, D, device="cuda")
        ex, _ = _scores(z, w, 1, False)
        tc, nres = _scores(z, w, 1, True)
        print(f"swap={swap} D={D} K={K} N={N}: equal={bool(torch.equal(ex, tc))} diff={int((ex != tc).sum())} resolved={nres}", flush=True)
lib.b200vq_tc_debug_desc_swap(0)

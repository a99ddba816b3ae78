"""Stand-alone launcher of the streaming linear kernels (profiling / timing aid, GPU box only)."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from video_tokenizer_b200 import _lib

ap = argparse.ArgumentParser()
ap.add_argument("--N", type=int, default=65536); ap.add_argument("--C", type=int, default=768)
ap.add_argument("--D", type=int, default=24); ap.add_argument("--reps", type=int, default=10)
a = ap.parse_args()
lib = _lib.load(); dev = torch.device("cuda:0"); torch.manual_seed(0)
N, C, D = a.N, a.C, a.D
x = torch.randn(N, C, device=dev); g = torch.randn(N, C, device=dev); y = torch.empty(N, C, device=dev)
w_in = torch.randn(D, C, device=dev) / C ** .5; b_in = torch.randn(D, device=dev)
w_out = torch.randn(C, D, device=dev) / D ** .5; b_out = torch.randn(C, device=dev)
z = torch.empty(N, D, device=dev); dw = torch.empty(C, D, device=dev); cs = torch.empty(C, device=dev)
flush = torch.empty(64 << 20, dtype=torch.float32, device=dev)
st = torch.cuda.current_stream().cuda_stream
fns = {
    "linear_in": lambda: lib.b200vq_linear_in(x.data_ptr(), C, N, C, w_in.data_ptr(), b_in.data_ptr(), D, z.data_ptr(), st),
    "linear_out": lambda: lib.b200vq_linear_out(z.data_ptr(), N, D, w_out.data_ptr(), 0, b_out.data_ptr(), C, y.data_ptr(), C, st),
    "weight_grad": lambda: lib.b200vq_weight_grad(g.data_ptr(), C, z.data_ptr(), N, C, D, dw.data_ptr(), 0, cs.data_ptr(), st),
}
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for name, fn in fns.items():
    for _ in range(3):
        assert fn() == 0
    tot = 0.0
    for _ in range(a.reps):
        flush.zero_()          # 256 MiB > L2
        e0.record(); assert fn() == 0; e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    ms = tot / a.reps
    print(f"{name}: {ms * 1e3:.1f} us  {N * C * 4 / ms / 1e6:.0f} GB/s of the N x C stream")
ref = torch.nn.functional.linear(x.double(), w_in.double(), b_in.double())
print("linear_in rel err vs fp64:", float((z.double() - ref).norm() / ref.norm()))

"""Tuning aid: time the streaming linear kernels alone (N=65536, C=768, D=24)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import video_tokenizer_b200 as V
from video_tokenizer_b200 import _lib
lib = V.load_library(); dev = torch.device("cuda")
N, C, D = 65536, 768, 24
xs = [torch.randn(N, C, device=dev) for _ in range(3)]     # rotate buffers: 3 x 201 MB > L2
ys = [torch.empty(N, C, device=dev) for _ in range(3)]
a = torch.randn(N, D, device=dev); w_in = torch.randn(D, C, device=dev) * 0.03; w_out = torch.randn(C, D, device=dev) * 0.1
b_in = torch.zeros(D, device=dev); b_out = torch.zeros(C, device=dev)
z = torch.empty(N, D, device=dev); dw = torch.empty(C, D, device=dev); cs = torch.empty(C, device=dev)
st = torch.cuda.current_stream().cuda_stream
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
i = [0]
def t(fn, reps=30):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3
def lin_in():
    i[0] += 1; _lib.check(lib.b200vq_linear_in(xs[i[0] % 3].data_ptr(), C, N, C, w_in.data_ptr(), b_in.data_ptr(), D, z.data_ptr(), st), "in")
def lin_out():
    i[0] += 1; _lib.check(lib.b200vq_linear_out(a.data_ptr(), N, D, w_out.data_ptr(), 0, b_out.data_ptr(), C, ys[i[0] % 3].data_ptr(), C, st), "out")
def lin_out_dc():
    i[0] += 1; _lib.check(lib.b200vq_linear_out(a.data_ptr(), N, D, w_in.data_ptr(), 1, None, C, ys[i[0] % 3].data_ptr(), C, st), "out")
def wgrad():
    i[0] += 1; _lib.check(lib.b200vq_weight_grad(xs[i[0] % 3].data_ptr(), C, a.data_ptr(), N, C, D, dw.data_ptr(), 0, cs.data_ptr(), st), "wg")
wsb = torch.empty(int(lib.b200vq_weight_grad_workspace_bytes(C, D)), dtype=torch.uint8, device=dev)
def wgrad_ws():
    i[0] += 1; _lib.check(lib.b200vq_weight_grad_ws(xs[i[0] % 3].data_ptr(), C, a.data_ptr(), N, C, D, dw.data_ptr(), 0, cs.data_ptr(), wsb.data_ptr(), wsb.numel(), st), "wgws")
cbk = torch.randn(8192, D, device=dev); didx = torch.randint(0, 8192, (N,), device=dev); pe = torch.randn(1024, C, device=dev)
def dec_proj():
    i[0] += 1; _lib.check(lib.b200vq_decode_project(didx.data_ptr(), N, cbk.data_ptr(), 8192, D, 1, w_out.data_ptr(), b_out.data_ptr(), C, pe.data_ptr(), 1024, ys[i[0] % 3].data_ptr(), C, None, st), "dp")
def dec_torch():
    i[0] += 1; q = torch.nn.functional.normalize(cbk[didx], dim=-1); torch.add(torch.nn.functional.linear(q, w_out, b_out).view(-1, 1024, C), pe, out=ys[i[0] % 3].view(-1, 1024, C))
ga = torch.empty(N, D, device=dev)
def grad_out():
    i[0] += 1; _lib.check(lib.b200vq_grad_out(xs[i[0] % 3].data_ptr(), C, a.data_ptr(), w_out.data_ptr(), N, C, D, None, ga.data_ptr(), dw.data_ptr(), cs.data_ptr(), wsb.data_ptr(), wsb.numel(), st), "go")
wt = w_out.t().contiguous()
def lin_in_tf32():
    i[0] += 1; _lib.check(lib.b200vq_linear_in(xs[i[0] % 3].data_ptr(), C, N, C, wt.data_ptr(), None, D, z.data_ptr(), st), "in")
mb = N * C * 4 / 1e6
for name, fn in (("linear_in(3xTF32)", lin_in), ("linear_out", lin_out), ("linear_out(dc)", lin_out_dc), ("weight_grad", wgrad), ("weight_grad_ws", wgrad_ws), ("grad_out (fused)", grad_out), ("decode_project+pe", dec_proj), ("decode torch eager", dec_torch)):
    us = t(fn); print(f"{name:20s} {us:7.1f} us  {mb / us:.2f} TB/s")

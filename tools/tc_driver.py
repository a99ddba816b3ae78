"""Stand-alone launcher of the scoring kernels (profiling / timing aid, GPU box only).

    python tools/tc_driver.py [--mode sample|argmin|argmax] [--N 65536] [--D 24] [--K 8192] [--reps 20] [--exact]
"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from video_tokenizer_b200 import _lib, functional as VF

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="sample")
ap.add_argument("--N", type=int, default=65536)
ap.add_argument("--D", type=int, default=24)
ap.add_argument("--K", type=int, default=8192)
ap.add_argument("--reps", type=int, default=20)
ap.add_argument("--exact", action="store_true")
a = ap.parse_args()
lib = _lib.load()
dev = torch.device("cuda:0")
torch.manual_seed(0)
mode = {"argmin": 0, "argmax": 1, "sample": 2}[a.mode]
z = torch.randn(a.N, a.D, device=dev)
w = torch.empty(a.K, a.D, device=dev); torch.nn.init.kaiming_uniform_(w)
emb, e2, _ = VF.prepare_codebook(w, True)
image = VF.prepare_codebook_tc(emb)
ws = torch.zeros(a.N * 4 + 64, dtype=torch.int32, device=dev)  # keys (2 words/row) + tc scratch
keys_ptr, second_ptr = ws.data_ptr(), ws.data_ptr() + a.N * 8
nres = torch.zeros(1, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream().cuda_stream

def run():
    ws.zero_()
    if a.exact:
        _lib.check(lib.b200vq_scores_partial(z.data_ptr(), 0, a.D, a.N, a.D, emb.data_ptr(), e2.data_ptr(), a.K, 0, mode, 1,
                                             1 / 0.03, 1, 4, 0, keys_ptr, st), "exact")
    else:
        _lib.check(lib.b200vq_scores_tc(z.data_ptr(), 0, a.D, a.N, a.D, image.data_ptr(), emb.data_ptr(), e2.data_ptr(),
                                        a.K, 0, mode, 1, 1 / 0.03, 1, 4, 0, keys_ptr, second_ptr, nres.data_ptr(), st), "tc")

for _ in range(3):
    run()
torch.cuda.synchronize()
nres.zero_()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.reps):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.reps
print(f"mode={a.mode} exact={a.exact} N={a.N} D={a.D} K={a.K} splits_env={os.environ.get('B200VQ_TC_SPLITS')}: "
      f"{ms * 1e3:.1f} us/call (incl. reset), {a.N / ms / 1e3:.1f} M tok/s, resolved/call={int(nres.item()) / a.reps:.1f}")

"""Tuning aid: time b200vq_scores_tc alone under the B200VQ_TC_DEBUG / B200VQ_TC_SPLITS knobs.
    python tools/tc_probe.py [D] [K] [N]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import video_tokenizer_b200 as V
from video_tokenizer_b200 import _lib, functional as VF

D = int(sys.argv[1]) if len(sys.argv) > 1 else 24
K = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
N = int(sys.argv[3]) if len(sys.argv) > 3 else 65536
lib = V.load_library()
dev = torch.device("cuda")
torch.manual_seed(0)
z = torch.randn(N, D, device=dev)
W = torch.empty(K, D, device=dev)
torch.nn.init.kaiming_uniform_(W)
emb, e2, _ = VF.prepare_codebook(W, True)
image = VF.prepare_codebook_tc(emb)
keys = torch.zeros(N, dtype=torch.int64, device=dev)
scratch = torch.zeros(2 * N + 4, dtype=torch.float32, device=dev)
st = torch.cuda.current_stream().cuda_stream
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


counter = torch.zeros(1, dtype=torch.int32, device=dev)


def run(mode):
    keys.zero_(); scratch.zero_()
    _lib.check(lib.b200vq_scores_tc(z.data_ptr(), 0, D, N, D, image.data_ptr(), emb.data_ptr(), e2.data_ptr(), K, 0,
                                    mode, 1, 1.0 / 0.03, 1, 4, 0, keys.data_ptr(), scratch.data_ptr(), counter.data_ptr(), st), "tc")


def timeit(mode, reps=20):
    for _ in range(3):
        run(mode)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        run(mode)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def overhead(reps=20):
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        keys.zero_(); scratch.zero_()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


print(f"D={D} K={K} N={N}  zeroing overhead {overhead():.1f} us")
counter.zero_(); run(_lib.MODE_SAMPLE); torch.cuda.synchronize()
print(f"sampler: {int(counter.item())} expanded groups = {int(counter.item()) / N:.1f} per row")
for dbg in sys.argv[4].split(",") if len(sys.argv) > 4 else ["0", "2", "3", "64", "6"]:
    for sp in (os.environ.get("PROBE_SPLITS", "0").split(",")):
        os.environ["B200VQ_TC_DEBUG"] = dbg
        if sp != "0":
            os.environ["B200VQ_TC_SPLITS"] = sp
        else:
            os.environ.pop("B200VQ_TC_SPLITS", None)
        print(f"debug={dbg:>3} splits={sp}: argmax {timeit(_lib.MODE_ARGMAX_COS):8.1f} us   sample {timeit(_lib.MODE_SAMPLE):8.1f} us",
              flush=True)

"""Tuning aid: time the second-generation sampler (b200vq_scores_tc, MODE_SAMPLE) under its knobs.
    python tools/ts_probe.py [D] [K] [N]
Knobs (environment, read per call): B200VQ_TS_SCOUT, B200VQ_TS_FLUSH_EARLY, B200VQ_TS_FLUSH_LATE,
B200VQ_TS_EARLY_ITERS, B200VQ_TC_DEBUG, B200VQ_TC_SPLITS.
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
counter = torch.zeros(1, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream().cuda_stream
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def run(mode=_lib.MODE_SAMPLE, zero=True):
    if zero:
        keys.zero_()
    _lib.check(lib.b200vq_scores_tc(z.data_ptr(), 0, D, N, D, image.data_ptr(), emb.data_ptr(), e2.data_ptr(), K, 0,
                                    mode, 1, 1.0 / 0.03, 1, 4, 0, keys.data_ptr(), scratch.data_ptr(), counter.data_ptr(), st), "tc")


def timeit(reps=20):
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        run(zero=False)     # keys only grow under atomicMax: same work every launch after the first
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def events():
    counter.zero_()
    run()
    torch.cuda.synchronize()
    return int(counter.item()) / N


def setenv(**kw):
    for k, v in kw.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = str(v)


print(f"D={D} K={K} N={N}")
# reference: exact kernel draw, for equality
ref = torch.zeros(N, dtype=torch.int64, device=dev)
_lib.check(lib.b200vq_scores_partial(z.data_ptr(), 0, D, N, D, emb.data_ptr(), e2.data_ptr(), K, 0, _lib.MODE_SAMPLE, 1,
                                     1.0 / 0.03, 1, 4, 0, ref.data_ptr(), st), "exact")
torch.cuda.synchronize()
for scout in os.environ.get("PROBE_SCOUT", "default,0,1,2,4,8").split(","):
    for fe, fl, ei in [tuple(int(v) for v in t.split(":")) for t in os.environ.get("PROBE_FLUSH", "8:32:2").split(",")]:
        for sp in os.environ.get("PROBE_SPLITS", "0").split(","):
            setenv(B200VQ_TS_SCOUT=None if scout == "default" else scout, B200VQ_TS_FLUSH_EARLY=fe, B200VQ_TS_FLUSH_LATE=fl,
                   B200VQ_TS_EARLY_ITERS=ei, B200VQ_TC_SPLITS=None if sp == "0" else sp, B200VQ_TC_DEBUG=None)
            ev = events()
            same = int((keys == ref).sum())
            t = timeit()
            line = f"scout={scout:>7} flush={fe}:{fl}:{ei} splits={sp}: {t:7.1f} us  {ev:5.1f} groups/row  equal-to-exact {same}/{N}"
            for dbg in os.environ.get("PROBE_DEBUG", "64,2,3").split(","):
                if dbg:
                    setenv(B200VQ_TC_DEBUG=dbg)
                    line += f"  dbg{dbg}={timeit():.1f}"
            print(line, flush=True)
setenv(B200VQ_TC_DEBUG=None, B200VQ_TS_SCOUT=None, B200VQ_TC_SPLITS=None)

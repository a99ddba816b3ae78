import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests.helpers import load_golden, golden_z
from tests.test_gpu_tc import _scores
g = load_golden("vq_edge_ties_zero")
z = golden_z(g).reshape(-1, 16).cuda(); w = torch.from_numpy(g["weight"]).cuda()
for mode in (0, 1):
    ex, _ = _scores(z, w, mode, False); tc, n = _scores(z, w, mode, True)
    print("mode", mode, "golden", g["idx"].flatten().tolist()); print(" exact", ex.tolist()); print(" tc   ", tc.tolist(), "resolved", n)

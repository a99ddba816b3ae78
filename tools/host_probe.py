"""Tuning aid: host (CPU) cost of one Bottleneck fwd+bwd step -- tiny batch so the GPU is never the bottleneck,
wall clock per step and a cProfile of 200 steps."""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import video_tokenizer_b200 as V

dev = torch.device("cuda")
spec = {"name": "bottleneck", "args": {"bottleneck_dim": 24, "norm": "none", "regularizer": {"name": "vq", "args": {
    "codebook_size": 8192, "l2_normalized": True, "stochastic": True, "stochastic_temperature": 0.03, "return_usage": True}}}}
m = V.make(spec, args={"token_nums": 1024, "input_dim": 768, "output_dim": 768}).to(dev).train()
x = torch.randn(1, 1024, 768, device=dev, requires_grad=True)
G = torch.randn(1, 1024, 768, device=dev)
gq = torch.tensor(0.1, device=dev)

def step():
    for p in m.parameters():
        p.grad = None
    x.grad = None
    out = m(x)
    torch.autograd.backward([out["output"], out["loss_q"]], [G, gq])

def fwd_only():
    with torch.no_grad():
        m(x)

for fn, name in ((step, "fwd+bwd"), (fwd_only, "fwd only (no_grad)")):
    for _ in range(20): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(300): fn()
    torch.cuda.synchronize(); print(f"{name}: {(time.perf_counter() - t0) / 300 * 1e6:.0f} us per step (host bound)")
pr = cProfile.Profile(); pr.enable()
for _ in range(200): step()
torch.cuda.synchronize(); pr.disable()
st = pstats.Stats(pr); st.sort_stats("tottime").print_stats(22)

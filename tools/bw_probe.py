"""Calibrate achievable HBM rates on the box: pure write (fill), pure read (sum), copy."""
import torch
dev = torch.device("cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def t(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e-3
for mb in (201, 805):
    n = mb * 1000 * 1000 // 4
    a = torch.empty(n, device=dev); b = torch.empty(n, device=dev)
    big = [torch.empty(n, device=dev) for _ in range(4)]
    i = [0]
    def fill():
        i[0] += 1; big[i[0] % 4].fill_(1.0)
    def read():
        i[0] += 1; big[i[0] % 4].sum()
    def copy():
        i[0] += 1; big[i[0] % 4].copy_(big[(i[0] + 1) % 4])
    print(f"{mb} MB: fill {n*4/t(fill)/1e12:.2f} TB/s  sum {n*4/t(read)/1e12:.2f} TB/s  copy(r+w) {2*n*4/t(copy)/1e12:.2f} TB/s")

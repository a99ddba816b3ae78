"""Timing of the 'sq' bottleneck configuration (reference models/larp_tokenizer.py:230-233: VectorQuantizer with the
frozen K = 196 560, D = 24 table, l2_norm=True) through the drop-in: fwd+bwd of 65 536 tokens, CUDA events."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import video_tokenizer_b200 as V

K, D, N = 196_560, 24, int(sys.argv[1]) if len(sys.argv) > 1 else 65536
torch.manual_seed(0)
q = V.VectorQuantizer(n_embed=K, embed_dim=D, l2_norm=True, beta=0.25, input_format="blc", predefined_codebook=None).cuda()
with torch.no_grad():
    q.embedding.weight.copy_(torch.nn.functional.normalize(torch.randn(K, D, device="cuda"), dim=-1))
q.embedding.weight.requires_grad = False
z = torch.randn(N // 1024, 1024, D, device="cuda", requires_grad=True)
G = torch.randn(N // 1024, 1024, D, device="cuda")
def step():
    out = q(z)
    ((out["output"] * G).sum() + out["loss_codebook"]).backward()
    z.grad = None
for _ in range(5): step()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(20): step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
print(json.dumps({"config": "sq: K=196560 D=24 l2 frozen, fwd+bwd", "tokens": N, "ms_per_step": ms, "tokens_per_s": N / ms * 1e3,
                  "issued_tflops": 2 * K * 80 * N / ms / 1e9, "reference_matrix_bytes_avoided": N * K * 4}))

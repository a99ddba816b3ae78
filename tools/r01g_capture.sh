# capture G: the entropy passes after the two-owner / split-stream rework (tests, probe, ncu of one fwd+bwd)
set -x
python -m pytest tests/test_entropy.py -m gpu -x -q 2>&1 | tail -3
python tools/entropy_probe.py gpurun_out/r01g_entropy_probe.jsonl 2>&1 | cut -c1-200
cat > /tmp/ent_one.py <<'PY'
import sys, torch
sys.path.insert(0, '.')
from video_tokenizer_b200 import functional as VF
torch.manual_seed(0)
z = torch.nn.functional.normalize(torch.randn(65536, 24, device='cuda'), dim=-1).requires_grad_(True)
e = torch.nn.functional.normalize(torch.randn(8192, 24, device='cuda'), dim=-1).requires_grad_(True)
for _ in range(2):
    z.grad = e.grad = None
    VF.entropy_loss_streamed(z, e, 0.01)[0].backward()
torch.cuda.synchronize()
PY
python /tmp/ent_one.py && ncu --set full --clock-control none --import-source on -k regex:entropy_pass -s 5 -c 5 -o gpurun_out/r01g_entropy python /tmp/ent_one.py > gpurun_out/ncu_ent.log 2>&1
ncu -i gpurun_out/r01g_entropy.ncu-rep --page details --csv > gpurun_out/r01g_entropy_details.csv 2>/dev/null
rm -f gpurun_out/r01g_entropy.ncu-rep
python -m pytest tests -m gpu -x -q 2>&1 | tail -3

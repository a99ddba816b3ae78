set -x
python bench.py > gpurun_out/r01f_bench_stochastic.json 2> gpurun_out/r01f_bench_stochastic.err
python bench.py --mode deterministic --no-cpu-baseline > gpurun_out/r01f_bench_deterministic.json 2>/dev/null
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r01f_bench_reference.json 2>/dev/null
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01f_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
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
python /tmp/ent_one.py && ncu --set full --clock-control none --import-source on -k regex:entropy_pass -s 5 -c 5 -o gpurun_out/r01f_entropy python /tmp/ent_one.py > gpurun_out/ncu_ent.log 2>&1
ncu -i gpurun_out/r01f_entropy.ncu-rep --page details --csv > gpurun_out/r01f_entropy_details.csv 2>/dev/null
tail -3 gpurun_out/ncu_ent.log
cat gpurun_out/r01f_bench_stochastic.json

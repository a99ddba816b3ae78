#!/bin/bash
# Round-2 profile capture (run under gpurun, 1 GPU).  Every ncu pass follows a plain run of the same command.
set -x
export PROBE_SCOUT=default PROBE_FLUSH=32:32:0 PROBE_DEBUG=64,2,3
python tools/ts_probe.py 24 8192 65536 > gpurun_out/r02_ts_probe.txt 2>&1
export PROBE_DEBUG=
python tools/ts_probe.py 24 8192 65536 > gpurun_out/r02_plain1.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:tc_sample -s 3 -c 1 -o gpurun_out/r02_tc_sample \
    python tools/ts_probe.py 24 8192 65536 > gpurun_out/r02_ncu1.log 2>&1
python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02_plain2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 120 --csv --log-file gpurun_out/r02_launches.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02_ncu2.log 2>&1
python tools/lin_probe.py > gpurun_out/r02_lin_probe.txt 2>&1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv > gpurun_out/r02_smi.txt

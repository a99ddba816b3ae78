# overlap-mask sweep (B200VQ_OVERLAP, see b200vq.cu: overlap_mask) at a given batch: bash tools/ov_probe.sh [batch] [masks...]
B=${1:-64}; shift
for m in ${@:-0 1 2 4 12 3 5 9 13 7 15}; do
  B200VQ_OVERLAP=$m python bench.py --no-extras --no-cpu-baseline --steps 40 --batch $B > /tmp/o.json 2>/dev/null
  python -c "
import json
d=json.loads(open('/tmp/o.json').read().strip().splitlines()[-1]); print('batch $B mask $m graph', round(d['ms_per_step'],4), 'eager', round(d['eager']['ms_per_step'],4))
"
done

for m in 13 5 9 7 13 0; do
  B200VQ_OVERLAP=$m python bench.py --no-extras --no-cpu-baseline --steps 40 > /tmp/o.json 2>/dev/null
  python -c "
import json
d=json.loads(open('/tmp/o.json').read().strip().splitlines()[-1]); print('mask $m graph', round(d['ms_per_step'],4), 'eager', round(d['eager']['ms_per_step'],4))
"
done

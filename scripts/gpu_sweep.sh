#!/bin/bash
# parity tests + sweep of one environment knob:  VAR=GSB_REFILL VALUES="1 4 8" bash scripts/gpu_sweep.sh
mkdir -p gpurun_out
[ -n "$NOTEST" ] || timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for v in $VALUES; do
  env $VAR=$v timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e $EXTRA 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$VAR=$v', round(d['value'],1), round(d['ms_per_step'],1), round(d.get('device_ms_per_step',0),1), {k: round(x,1) for k,x in d['stage_ms_per_step'].items()})
    else: print(l.rstrip()[:200])
"
done

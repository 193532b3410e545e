#!/bin/bash
# parity tests + sweep over environment settings:  CONFIGS="GSB_VOTE=0,GSB_REFILL=24 GSB_VOTE=12,GSB_REFILL=16" EXTRA="--scene cornell" bash scripts/gpu_sweep2.sh
mkdir -p gpurun_out
[ -n "$NOTEST" ] || timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for cfg in $CONFIGS; do
  env $(echo $cfg | tr ',' ' ') timeout 300 python bench.py --steps ${STEPS:-3} --warmup 3 --spp-per-step ${SPP:-64} --no-cpu-baseline --no-e2e --no-instrumented $EXTRA 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$cfg', round(d['value'],1), 'ms', round(d['ms_per_step'],1), {k: round(x,1) for k,x in d['stage_ms_per_step'].items()})
    else: print(l.rstrip()[:200])
"
done

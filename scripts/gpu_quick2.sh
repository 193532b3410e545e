#!/bin/bash
# quick pass: parity tests, roofline denominators, short benches of the three scenes
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q -x 2>&1 | tail -6
timeout 120 gpuspectral_b200/lib/gsb_peaks > gpurun_out/peaks.json; cat gpurun_out/peaks.json
for sc in cornell gallery field; do
  if [ $sc = field ]; then A="--steps 3 --spp-per-step 16"; elif [ $sc = gallery ]; then A="--steps 3 --spp-per-step 32"; else A="--steps 4 --spp-per-step 128"; fi
  timeout 400 python bench.py --scene $sc $A --no-e2e --no-cpu-baseline --no-instrumented 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$sc', round(d['value'],1), 'ms', round(d['ms_per_step'],1), {k: round(x,1) for k,x in d['stage_ms_per_step'].items()})
    else: print(l.rstrip()[:200])
"
done

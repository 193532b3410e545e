#!/bin/bash
# A/B of compiler flags for the wavefront kernels: rebuilds on the box per variant, runs the parity tests and short benches
mkdir -p gpurun_out
IFS='|'
for v in $VARIANTS; do
  unset IFS
  make -C gpuspectral_b200 clean > /dev/null
  make -C gpuspectral_b200 WF_FLAGS="$v" > /dev/null 2>&1 || { echo "[$v]: build failed"; continue; }
  echo "== WF_FLAGS=[$v]"
  timeout 400 python -m pytest tests -m gpu -q 2>&1 | tail -4
  for sc in cornell gallery; do
    if [ $sc = gallery ]; then A="--steps 3 --spp-per-step 32"; else A="--steps 3 --spp-per-step 64"; fi
    timeout 400 python bench.py --scene $sc $A --no-e2e --no-cpu-baseline --no-instrumented 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$sc', round(d['value'],1), 'ms', round(d['ms_per_step'],1), {k: round(x,1) for k,x in d['stage_ms_per_step'].items()})
    else: print(l.rstrip()[:200])
"
  done
  IFS='|'
done

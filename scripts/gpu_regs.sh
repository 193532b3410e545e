#!/bin/bash
# register-budget A/B of the traversal / shading kernels: rebuilds the core on the box per variant (VARIANTS="TRACE=7 TRACE=9 SHADE=6")
mkdir -p gpurun_out
for v in $VARIANTS; do
  k=${v%%=*}; n=${v##*=}
  make -C gpuspectral_b200 clean > /dev/null
  if [ $k = SMEM_STACK ]; then D=-DGSB_SMEM_STACK=$n; elif [ $k = WORK_CHUNK ]; then D=-DGSB_WORK_CHUNK=$n; else D=-DGSB_${k}_MIN_BLOCKS=$n; fi
  make -C gpuspectral_b200 EXTRA_DEFS=$D > /dev/null 2>&1 || { echo "$v: build failed"; continue; }
  grep -A3 "k_extendILb0\|k_shadeILi0" gpuspectral_b200/build/wavefront.ptxas.log | grep "Used\|spill" | tr '\n' ' '; echo
  for sc in cornell ${SCENES}; do
    if [ $sc = field ]; then A="--steps 3 --spp-per-step 16"; elif [ $sc = gallery ]; then A="--steps 3 --spp-per-step 32"; else A="--steps 3 --spp-per-step 64"; fi
    timeout 400 python bench.py --scene $sc $A --no-e2e --no-cpu-baseline --no-instrumented 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$v $sc', round(d['value'],1), 'ms', round(d['ms_per_step'],1), {k: round(x,1) for k,x in d['stage_ms_per_step'].items()})
    else: print(l.rstrip()[:200])
"
  done
done

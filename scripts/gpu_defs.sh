#!/bin/bash
# A/B of compile-time variants: rebuilds the core on the box per entry of VARIANTS (";"-separated nvcc -D lists, "base" = none)
mkdir -p gpurun_out
IFS=';' read -ra VS <<< "$VARIANTS"
for v in "${VS[@]}"; do
  make -C gpuspectral_b200 clean > /dev/null
  if [ "$v" = base ]; then D=""; else D="$v"; fi
  make -C gpuspectral_b200 -j8 EXTRA_DEFS="$D" > /dev/null 2>&1 || { echo "$v: build failed"; continue; }
  grep -A3 "k_extendILb0\|k_shadeILi0" gpuspectral_b200/build/wavefront.ptxas.log | grep "Used\|spill" | tr '\n' ' '; echo
  for sc in cornell ${SCENES}; do
    if [ $sc = field ]; then A="--steps 3 --spp-per-step 16"; elif [ $sc = gallery ]; then A="--steps 3 --spp-per-step 32"; else A="--steps 4 --spp-per-step 128"; fi
    timeout 400 python bench.py --scene $sc $A --no-e2e --no-cpu-baseline --no-instrumented --no-extra 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('[$v] $sc', round(d['value'],1), 'ms', round(d['ms_per_step'],1), {k: round(x,1) for k,x in d['stage_ms_per_step'].items()})
    else: print(l.rstrip()[:200])
"
  done
done

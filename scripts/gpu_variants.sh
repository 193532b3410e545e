#!/bin/bash
# On the GPU box: the bench scenes (SCENES, default "cornell") with every prebuilt variant of the core under gpurun_variants/
# (scripts/build_variant.sh) and with the tree's own build ("base"), REPS times each, interleaved.
mkdir -p gpurun_out
cp gpuspectral_b200/lib/libgpuspectral_b200.so /tmp/base_core.so
for rep in $(seq 1 ${REPS:-2}); do
for v in base $(ls gpurun_variants 2>/dev/null); do
  if [ $v = base ]; then cp /tmp/base_core.so gpuspectral_b200/lib/libgpuspectral_b200.so; else cp gpurun_variants/$v/libgpuspectral_b200.so gpuspectral_b200/lib/libgpuspectral_b200.so; fi
  for sc in ${SCENES:-cornell}; do
    if [ $sc = field ]; then A="--steps 3 --spp-per-step 16"; elif [ $sc = gallery ]; then A="--steps 3 --spp-per-step 64"; else A="--steps 4 --spp-per-step 128"; fi
    timeout 400 python bench.py --scene $sc $A --no-e2e --no-cpu-baseline --no-instrumented --no-extra 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('[$v] $sc', round(d['value'],1), 'ms', round(d['ms_per_step'],1), {k: round(x,1) for k,x in d['stage_ms_per_step'].items()})
"
  done
done
done
cp /tmp/base_core.so gpuspectral_b200/lib/libgpuspectral_b200.so

#!/bin/bash
# treelet refinement A/B: parity tests with the refinement on, then gallery / field benches per number of passes
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log
for p in ${PASSES:-0 1 2 3}; do
for sc in gallery field; do
  if [ $sc = field ]; then A="--steps 3 --spp-per-step 16"; else A="--steps 3 --spp-per-step 32"; fi
  GSB_TREELET_PASSES=$p timeout 600 python bench.py --scene $sc $A --no-e2e --no-cpu-baseline 2> gpurun_out/tl_${sc}_$p.err > gpurun_out/tl_${sc}_$p.json
  tail -3 gpurun_out/tl_${sc}_$p.err
  python - gpurun_out/tl_${sc}_$p.json $p $sc <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    b = d['bvh']
    print(sys.argv[3], 'passes', sys.argv[2], 'Mpaths/s', round(d['value'], 1), 'stages', {k: round(v, 1) for k, v in d['stage_ms_per_step'].items()}, 'B/ray', round(d['roofline']['bytes_per_ray']),
          'fetch', {k: round(v, 2) for k, v in d['roofline']['fetch_per_ray'].items()}, 'build_ms', round(b['build_ms'], 1), 'sah', round(b['sah_cost_before'], 1), '->', round(b['sah_cost_after'], 1), 'nodes', b['blas_nodes'])
except Exception as e:
    print(sys.argv[1], 'unreadable', e)
PY
done; done

#!/bin/bash
# r1h / r1i pass: parity tests, smoke, benches of the three scenes (+ TLAS refinement A/B on the field), launch list,
# full ncu of the stage kernels (cornell and field)
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
summ() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], 'value', round(d['value'], 1), 'Mrays', round(d['mrays_per_s']), 'ms/step', round(d['ms_per_step'], 1), 'e2e', d['e2e'] and round(d['e2e']['value'], 1))
    b = d.get('bvh') or {}
    print('  stages', {k: round(v, 1) for k, v in d['stage_ms_per_step'].items()}, 'frac', round(d['roofline']['frac'], 3), 'B/ray', round(d['roofline']['bytes_per_ray']),
          'fetch', {k: round(v, 2) for k, v in d['roofline']['fetch_per_ray'].items()},
          'build_ms', round(b.get('build_ms', 0), 1), 'sah', round(b.get('sah_cost_before', 0), 1), '->', round(b.get('sah_cost_after', 0), 1))
except Exception as e:
    print(sys.argv[1], 'unreadable', e)
PY
}
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -3 gpurun_out/bench.err; summ gpurun_out/bench.json
timeout 600 python bench.py --scene gallery --steps 4 --spp-per-step 64 --no-e2e > gpurun_out/bench_gallery.json 2> gpurun_out/bench_gallery.err; tail -3 gpurun_out/bench_gallery.err; summ gpurun_out/bench_gallery.json
timeout 600 python bench.py --scene field --steps 3 --spp-per-step 16 --no-e2e --no-cpu-baseline > gpurun_out/bench_field.json 2> gpurun_out/bench_field.err; tail -3 gpurun_out/bench_field.err; summ gpurun_out/bench_field.json
GSB_TREELET_PASSES_TLAS=2 timeout 600 python bench.py --scene field --steps 3 --spp-per-step 16 --no-e2e --no-cpu-baseline > gpurun_out/bench_field_tlas2.json 2> gpurun_out/bench_field_tlas2.err; summ gpurun_out/bench_field_tlas2.json
SMALL="python bench.py --steps 1 --warmup 1 --spp-per-step 8 --no-cpu-baseline --no-e2e --no-instrumented"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade|k_shadow" -s 30 -c 3 -o gpurun_out/prof_cornell $SMALL > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade|k_shadow" -s 30 -c 5 -o gpurun_out/prof_field $SMALL --scene field > gpurun_out/ncu_full_field.log 2>&1
tail -2 gpurun_out/ncu_full_field.log
ls -la gpurun_out | head -40

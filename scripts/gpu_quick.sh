#!/bin/bash
# quick GPU pass: parity tests, smoke, bench (compact output), optional ncu of the stage kernels
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py $BENCH_ARGS > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
print('value', round(d['value'], 1), 'Mrays', round(d['mrays_per_s']), 'ms/step', round(d['ms_per_step'], 1), 'device', round(d.get('device_ms_per_step', 0), 1), 'e2e', d['e2e'] and round(d['e2e']['value'], 1))
print('stages', {k: round(v, 1) for k, v in d['stage_ms_per_step'].items()}, 'frac', round(d['roofline']['frac'], 3), 'launches', d['gpu_launches'])
PY
if [ -n "$NCU" ]; then
SMALL="python bench.py --steps 1 --warmup 1 --spp-per-step 4 --no-cpu-baseline --no-e2e --no-instrumented"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade|k_shadow" -s 30 -c 6 -o gpurun_out/prof $SMALL > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
fi

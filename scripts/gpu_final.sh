#!/bin/bash
# closing pass of a round: the GPU suite, smoke, the default bench line and the reference arm exactly as the driver runs them
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q 2>&1 | tail -8 > gpurun_out/pytest_gpu_all.log; cat gpurun_out/pytest_gpu_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -c 900 gpurun_out/bench_reference.json
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/bench.json').read().strip().splitlines()[-1])
print('value', round(d['value'], 1), 'e2e', round(d['e2e']['value'], 1), 'cpu', d['cpu_baseline'], 'clocks', d['clocks'])
PY

#!/bin/bash
# A/B of one build: GPU suite, then the three bench scenes (no e2e / CPU legs)
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/pytest_gpu_all.log; cat gpurun_out/pytest_gpu_all.log
summ() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], 'value', round(d['value'], 1), 'ms/step', round(d['ms_per_step'], 1), 'stages', {k: round(v, 1) for k, v in d['stage_ms_per_step'].items()})
except Exception as e:
    print(sys.argv[1], 'unreadable', e)
PY
}
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -2 gpurun_out/bench.err; summ gpurun_out/bench.json
timeout 600 python bench.py --scene gallery --steps 4 --spp-per-step 64 --no-e2e --no-cpu-baseline > gpurun_out/bench_gallery.json 2> gpurun_out/bench_gallery.err; summ gpurun_out/bench_gallery.json
timeout 600 python bench.py --scene field --steps 3 --spp-per-step 16 --no-e2e --no-cpu-baseline > gpurun_out/bench_field.json 2> gpurun_out/bench_field.err; summ gpurun_out/bench_field.json

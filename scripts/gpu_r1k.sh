#!/bin/bash
# r1k pass (final build of round 1): GPU suite, smoke, reference arm, the three bench scenes, launch list, full ncu of the stage kernels
mkdir -p gpurun_out
bash scripts/gpu_final.sh
summ() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], 'value', round(d['value'], 1), 'Mrays', round(d['mrays_per_s']), 'ms/step', round(d['ms_per_step'], 1), 'e2e', d['e2e'] and round(d['e2e']['value'], 1))
    print('  stages', {k: round(v, 1) for k, v in d['stage_ms_per_step'].items()}, 'frac', round(d['roofline']['frac'], 3), 'l2frac', round(d['roofline'].get('frac_of_l2_peak', 0), 3))
except Exception as e:
    print(sys.argv[1], 'unreadable', e)
PY
}
summ gpurun_out/bench.json
timeout 600 python bench.py --scene gallery --steps 4 --spp-per-step 64 --no-e2e > gpurun_out/bench_gallery.json 2> gpurun_out/bench_gallery.err; summ gpurun_out/bench_gallery.json
timeout 600 python bench.py --scene field --steps 3 --spp-per-step 16 --no-e2e --no-cpu-baseline > gpurun_out/bench_field.json 2> gpurun_out/bench_field.err; summ gpurun_out/bench_field.json
bash scripts/gpu_ncu2.sh
ls -la gpurun_out | head -40

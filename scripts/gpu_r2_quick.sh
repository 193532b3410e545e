#!/bin/bash
# quick r2 loop: GPU suite + the three bench scenes (no cpu baseline, no e2e); TAG names the output files
TAG=${TAG:-r2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/${TAG}_pytest.log; cat gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/${TAG}_cornell.json 2> gpurun_out/${TAG}_cornell.err || tail -5 gpurun_out/${TAG}_cornell.err
timeout 600 python bench.py --scene gallery --steps 3 --spp-per-step 64 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_gallery.json 2> gpurun_out/${TAG}_gallery.err || tail -5 gpurun_out/${TAG}_gallery.err
timeout 600 python bench.py --scene field --steps 3 --spp-per-step 16 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_field.json 2> gpurun_out/${TAG}_field.err || tail -5 gpurun_out/${TAG}_field.err
python - <<PY
import json
for n in ['cornell','gallery','field']:
    try:
        d=json.loads(open(f'gpurun_out/${TAG}_{n}.json').read().strip().splitlines()[-1])
        r=d.get('roofline',{})
        print(n, 'Mpaths/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],1), {k: round(v,1) for k,v in d.get('stage_ms_per_step',{}).items()}, 'fetch', r.get('fetch_per_ray'), 'frac', r.get('frac'))
    except Exception as e: print(n, 'err', e)
PY

#!/bin/bash
# A/B over environment settings: each argument is "TAG:ENV=VAL,ENV2=VAL2" (or "TAG:" for defaults); three bench scenes each
mkdir -p gpurun_out
for spec in "$@"; do
  tag=${spec%%:*}; envs=${spec#*:}
  (
    IFS=','; for kv in $envs; do [ -n "$kv" ] && export "$kv"; done; unset IFS
    timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_cornell.json 2> gpurun_out/${tag}_cornell.err || tail -5 gpurun_out/${tag}_cornell.err
    timeout 600 python bench.py --scene gallery --steps 3 --spp-per-step 64 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_gallery.json 2> gpurun_out/${tag}_gallery.err || tail -5 gpurun_out/${tag}_gallery.err
    timeout 600 python bench.py --scene field --steps 3 --spp-per-step 16 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_field.json 2> gpurun_out/${tag}_field.err || tail -5 gpurun_out/${tag}_field.err
  )
  python - <<PY
import json
for n in ['cornell','gallery','field']:
    try:
        d=json.loads(open(f'gpurun_out/${tag}_{n}.json').read().strip().splitlines()[-1])
        r=d.get('roofline',{})
        f=r.get('fetch_per_ray') or {}
        print('${tag}', n, 'Mpaths/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],1), {k: round(v,1) for k,v in d.get('stage_ms_per_step',{}).items()}, 'fetch', {k: round(v,2) for k,v in f.items()})
    except Exception as e: print('${tag}', n, 'err', e)
PY
done

#!/bin/bash
# r2 full pass: GPU suite (with timings), smoke, the default bench line (extras, e2e, cpu baseline) and the reference arm
TAG=${TAG:-r2}
mkdir -p gpurun_out
rm -f gpurun_out/parity_full_size.jsonl
timeout 1500 python -m pytest tests -m gpu -q -x --durations=12 2>&1 | tail -30 > gpurun_out/${TAG}_pytest.log; cat gpurun_out/${TAG}_pytest.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || tail -5 gpurun_out/${TAG}_bench.err
timeout 600 python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err || tail -5 gpurun_out/${TAG}_bench_reference.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_bench.json').read().strip().splitlines()[-1])
print('cornell', round(d['value'],1), 'e2e', round(d['e2e']['value'],1) if d.get('e2e') else None, 'frac', d['roofline']['bound'], round(d['roofline']['frac'],3), 'shade', {k:(round(v,3) if isinstance(v,float) else v) for k,v in d['roofline_shade'].items() if k in ('achieved_gbs','frac_hbm','vertices_per_launch')}, 'cpu', d.get('cpu_baseline'))
for k,v in (d.get('extra_workloads') or {}).items():
    print(k, round(v['value'],1) if 'value' in v else v, v.get('stage_ms_per_step'), v.get('roofline',{}).get('frac'))
r=json.loads(open('gpurun_out/${TAG}_bench_reference.json').read().strip().splitlines()[-1])
print('reference', r['value'], r['cpu_baseline'])
PY
cat gpurun_out/parity_full_size.jsonl

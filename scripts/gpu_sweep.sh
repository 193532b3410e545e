#!/bin/bash
# parity tests + slot-count sweep of the work-stealing wavefront
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for s in 262144 524288 1048576 2097152 4194304; do
  echo "GSB_SLOTS=$s"
  GSB_SLOTS=$s python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['value'], d['mrays_per_s'], d['ms_per_step'], d['stage_ms_per_step'], d['roofline']['launches'], d['roofline']['frac'])
    else: print(l.rstrip())
"
done

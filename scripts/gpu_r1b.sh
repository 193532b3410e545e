#!/bin/bash
# second GPU pass (work-stealing wavefront): parity tests, smoke, bench, slot sweep, launch list, ncu captures
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/pytest_gpu.log; tail -15 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -5 gpurun_out/bench.err; cat gpurun_out/bench.json
for s in 262144 524288 2097152 4194304; do
  echo "GSB_SLOTS=$s"
  GSB_SLOTS=$s python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tee gpurun_out/sweep_$s.json | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(d['value'], d['mrays_per_s'], d['ms_per_step'], d['stage_ms_per_step'], d['roofline']['launches'], d['roofline']['frac'])
    else: print(l.rstrip())
"
done
SMALL="python bench.py --steps 1 --warmup 1 --spp-per-step 4 --no-cpu-baseline --no-e2e"
$SMALL > gpurun_out/small_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade|k_shadow|k_finish" -s 40 -c 8 -o gpurun_out/prof_r1b $SMALL > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out

#!/bin/bash
# first GPU pass: parity tests, smoke, bench, launch list, one full ncu capture of the traversal kernel
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv
python -m pytest tests -m gpu -x -q 2>&1 | tail -30 > gpurun_out/pytest_gpu.log; tail -15 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -5 gpurun_out/bench.err; cat gpurun_out/bench.json
SMALL="python bench.py --steps 1 --warmup 1 --spp-per-step 4 --no-cpu-baseline --no-e2e"
$SMALL > gpurun_out/small_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1
$SMALL > gpurun_out/small_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_extend -s 12 -c 2 -o gpurun_out/prof_extend $SMALL > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out

#!/bin/bash
# profiler-only pass: launch list + full capture of the production stage kernels (no instrumented pass)
mkdir -p gpurun_out
SMALL="python bench.py --steps 1 --warmup 1 --spp-per-step 4 --no-cpu-baseline --no-e2e --no-instrumented"
$SMALL > gpurun_out/small_plain.log 2>&1 || { tail -5 gpurun_out/small_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade|k_shadow" -s ${SKIP:-30} -c ${COUNT:-6} -o gpurun_out/prof $SMALL > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log

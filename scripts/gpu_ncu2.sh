#!/bin/bash
# profiler pass for the 8 M-slot default: 32 spp per step = four full waves, so launch 30.. is a steady-state iteration
mkdir -p gpurun_out
SMALL="python bench.py --steps 1 --warmup 1 --spp-per-step 32 --no-cpu-baseline --no-e2e --no-instrumented"
$SMALL > gpurun_out/small_plain.log 2>&1 || { tail -5 gpurun_out/small_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade|k_shadow" -s 30 -c 3 -o gpurun_out/prof_cornell $SMALL > gpurun_out/ncu_full.log 2>&1
tail -1 gpurun_out/ncu_full.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade|k_shadow" -s 30 -c 5 -o gpurun_out/prof_field $SMALL --scene field > gpurun_out/ncu_full_field.log 2>&1
tail -1 gpurun_out/ncu_full_field.log

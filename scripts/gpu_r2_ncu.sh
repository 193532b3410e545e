#!/bin/bash
# profiler pass: launch list + full capture of steady-state stage kernels.  SCENE=cornell|gallery|field, TAG names outputs
TAG=${TAG:-r2}; SCENE=${SCENE:-cornell}
mkdir -p gpurun_out
SMALL="python bench.py --scene $SCENE --steps 1 --warmup 1 --spp-per-step ${SPP:-32} --no-cpu-baseline --no-e2e --no-instrumented"
$SMALL > gpurun_out/${TAG}_small_plain.log 2>&1 || { tail -5 gpurun_out/${TAG}_small_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/${TAG}_${SCENE}_launches.csv $SMALL > gpurun_out/${TAG}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_extend|k_shade|k_shadow" -s ${SKIP:-30} -c ${COUNT:-3} -o gpurun_out/${TAG}_prof_${SCENE} $SMALL > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -1 gpurun_out/${TAG}_ncu_full.log

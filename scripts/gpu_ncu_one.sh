#!/bin/bash
# one steady-state launch of one kernel under ncu --set full: KERNEL (regex), SKIP (launches of it to pass over), SCENE, SPP, TAG
TAG=${TAG:-r2}; SCENE=${SCENE:-cornell}
mkdir -p gpurun_out
SMALL="python bench.py --scene $SCENE --steps 1 --warmup 1 --spp-per-step ${SPP:-128} --no-cpu-baseline --no-e2e --no-instrumented --no-extra"
GSB_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${KERNEL:-k_regen}" -s ${SKIP:-300} -c ${COUNT:-1} -o gpurun_out/${TAG}_one_${SCENE} $SMALL > gpurun_out/${TAG}_ncu_one.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_one.log

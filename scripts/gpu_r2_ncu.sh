#!/bin/bash
# profiler pass: launch list + full capture of steady-state stage kernels + FP32 op counts of k_shade.
# SCENE=cornell|gallery|field, TAG names outputs
TAG=${TAG:-r2}; SCENE=${SCENE:-cornell}
mkdir -p gpurun_out
SMALL="python bench.py --scene $SCENE --steps 1 --warmup 1 --spp-per-step ${SPP:-32} --no-cpu-baseline --no-e2e --no-instrumented --no-extra"
GSB_GRAPH=0 $SMALL > gpurun_out/${TAG}_small_plain.log 2>&1 || { tail -5 gpurun_out/${TAG}_small_plain.log; exit 1; }
GSB_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_${SCENE}_launches.csv $SMALL > gpurun_out/${TAG}_ncu_launches.log 2>&1
GSB_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_regen|k_extend|k_shade|k_shadow" -s ${SKIP:-40} -c ${COUNT:-4} -o gpurun_out/${TAG}_prof_${SCENE} $SMALL > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -1 gpurun_out/${TAG}_ncu_full.log
GSB_GRAPH=0 timeout 900 ncu --clock-control none -k regex:"k_shade" -s ${SKIP:-10} -c 2 --csv --metrics smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fp32_pred_on.sum,smsp__inst_executed_pipe_xu.sum,smsp__thread_inst_executed.sum,gpu__time_duration.sum --log-file gpurun_out/${TAG}_${SCENE}_shade_ops.csv $SMALL > gpurun_out/${TAG}_ncu_ops.log 2>&1
tail -3 gpurun_out/${TAG}_${SCENE}_shade_ops.csv

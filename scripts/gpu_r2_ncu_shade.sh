mkdir -p gpurun_out
SMALL="python bench.py --scene cornell --steps 1 --warmup 1 --spp-per-step 32 --no-cpu-baseline --no-e2e --no-instrumented --no-extra"
GSB_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_shade" -s 12 -c 1 -o gpurun_out/r2i_prof_shade $SMALL > gpurun_out/r2i_ncu_shade.log 2>&1
tail -1 gpurun_out/r2i_ncu_shade.log

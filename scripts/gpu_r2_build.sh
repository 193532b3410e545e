#!/bin/bash
# launch list of the field scene's BVH builds (three uploads): which kernels the build time goes to
TAG=${TAG:-r2u}
mkdir -p gpurun_out
python scripts/gpu_build_probe.py > gpurun_out/${TAG}_build_probe.log 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_field_build_launches.csv \
    python scripts/gpu_build_probe.py > gpurun_out/${TAG}_build_ncu.log 2>&1
cat gpurun_out/${TAG}_build_probe.log

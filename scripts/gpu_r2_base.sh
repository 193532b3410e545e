#!/bin/bash
# r2 baseline: GPU suite after the oracle pin + the three bench scenes on this pool's box (round-1 kernels)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/r2_pytest_base.log; cat gpurun_out/r2_pytest_base.log
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/r2_base_cornell.json 2> gpurun_out/r2_base_cornell.err; tail -c 600 gpurun_out/r2_base_cornell.json
timeout 600 python bench.py --scene gallery --steps 3 --spp-per-step 64 --no-e2e --no-cpu-baseline > gpurun_out/r2_base_gallery.json 2> gpurun_out/r2_base_gallery.err; tail -c 300 gpurun_out/r2_base_gallery.json
timeout 600 python bench.py --scene field --steps 3 --spp-per-step 16 --no-e2e --no-cpu-baseline > gpurun_out/r2_base_field.json 2> gpurun_out/r2_base_field.err; tail -c 300 gpurun_out/r2_base_field.json

#!/bin/bash
# multi-GPU pass (gpurun --gpus N): the contract launch of bench.py under torchrun, then BASELINE configs[4] (4096x2160, sample-sharded)
N=${N:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
tail -3 gpurun_out/bench_n$N.err; tail -1 gpurun_out/bench_n$N.json | cut -c1-600
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 4 --warmup 3 --width 4096 --height 2160 --spp-per-step 32 --no-cpu-baseline --no-e2e > gpurun_out/bench_4k_n$N.json 2> gpurun_out/bench_4k_n$N.err
tail -3 gpurun_out/bench_4k_n$N.err; tail -1 gpurun_out/bench_4k_n$N.json | cut -c1-600

#!/bin/bash
# multi-GPU pass (gpurun --gpus N): the 2-rank NCCL sharding test, the contract launch of bench.py under torchrun, then
# BASELINE configs[4] (4096x2160, sample-sharded), then the reference arm the way the driver launches it
N=${N:-2}; TAG=${TAG:-r2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 600 python -m pytest tests/test_gpu_sharding.py -q -x 2>&1 | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 8 --warmup 3 > gpurun_out/${TAG}_bench_cornell_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
tail -3 gpurun_out/${TAG}_bench_n$N.err; tail -1 gpurun_out/${TAG}_bench_cornell_n$N.json | cut -c1-400
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 4 --warmup 3 --width 4096 --height 2160 --spp-per-step 32 --no-e2e > gpurun_out/${TAG}_bench_cornell_4k_n$N.json 2> gpurun_out/${TAG}_bench_4k_n$N.err
tail -3 gpurun_out/${TAG}_bench_4k_n$N.err; tail -1 gpurun_out/${TAG}_bench_cornell_4k_n$N.json | cut -c1-400
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 4 --warmup 1 > gpurun_out/${TAG}_bench_reference_n$N.json 2> gpurun_out/${TAG}_bench_reference_n$N.err
tail -1 gpurun_out/${TAG}_bench_reference_n$N.json | cut -c1-300

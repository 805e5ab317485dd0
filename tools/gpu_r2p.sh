#!/bin/bash
# round 2, GPU call P (8 GPUs): final bench line at N = 8 with the e2e breakdown
set -x
mkdir -p gpurun_out
NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r2p_bench_n8.json 2> gpurun_out/r2p_bench_n8.err
tail -c 2500 gpurun_out/r2p_bench_n8.json
NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r2p_bench_n4.json 2> gpurun_out/r2p_bench_n4.err
tail -c 1200 gpurun_out/r2p_bench_n4.json

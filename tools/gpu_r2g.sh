#!/bin/bash
# round 2, GPU call G (N GPUs): NCCL C entry test + bench at N ranks.  usage: bash tools/gpu_r2g.sh N
N=${1:-2}
set -x
mkdir -p gpurun_out
nvidia-smi -L | head -8
if [ "$N" = "2" ]; then
  timeout 600 python -m pytest tests/test_shard.py -x -q -m gpu > gpurun_out/r2g_pytest_shard_n2.log 2>&1
  tail -5 gpurun_out/r2g_pytest_shard_n2.log
fi
NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2g_bench_n$N.json 2> gpurun_out/r2g_bench_n$N.err
tail -c 1500 gpurun_out/r2g_bench_n$N.json
tail -5 gpurun_out/r2g_bench_n$N.err

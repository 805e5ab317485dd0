#!/bin/bash
# round 2, GPU call B: bulk-copy kernel parity + timing, remaining vendor comparison, NCCL C entry at world 1
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stream_path.py -x -q -m gpu > gpurun_out/r2b_pytest_stream.log 2>&1
tail -5 gpurun_out/r2b_pytest_stream.log
timeout 600 python tools/r2_csr_experiments.py --out gpurun_out/r2b_csr_experiments.jsonl > gpurun_out/r2b_csr_experiments.log 2>&1
tail -3 gpurun_out/r2b_csr_experiments.log
timeout 300 python -m pytest tests/test_shard.py tests/test_gpu_vendor_parity.py -x -q -m gpu > gpurun_out/r2b_pytest_shard_vendor.log 2>&1
tail -5 gpurun_out/r2b_pytest_shard_vendor.log
timeout 900 python baseline/compare.py --configs c3,c4,c5 --out gpurun_out/r2b_vendor.jsonl > gpurun_out/r2b_vendor.log 2>&1
tail -3 gpurun_out/r2b_vendor.log

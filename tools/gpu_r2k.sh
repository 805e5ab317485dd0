#!/bin/bash
# round 2, GPU call K: pipelined lean kernel -- parity, C2 timings, C5 bf16 timings
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stream_path.py -x -q -m gpu > gpurun_out/r2k_pytest_stream.log 2>&1
tail -4 gpurun_out/r2k_pytest_stream.log
timeout 600 python tools/r2_c2_graph.py gpurun_out/r2k_c2_graph.jsonl > gpurun_out/r2k_c2_graph.log 2>&1
tail -2 gpurun_out/r2k_c2_graph.log | cut -c1-300
timeout 600 python tools/r2_csr_experiments.py --policies 0 --out gpurun_out/r2k_csr_experiments.jsonl > gpurun_out/r2k_csr_experiments.log 2>&1
tail -2 gpurun_out/r2k_csr_experiments.log | cut -c1-300

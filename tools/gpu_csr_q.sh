#!/bin/bash
# CSR parity + timings after the inline segment queue change
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/time_kernels.py --cases c2,c2l,c4csr,c5,c5l > gpurun_out/time_csr_q.jsonl 2> gpurun_out/time_csr_q.err; echo "time rc=$?"; cut -c1-200 gpurun_out/time_csr_q.jsonl; tail -3 gpurun_out/time_csr_q.err

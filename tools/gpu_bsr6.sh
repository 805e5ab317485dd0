#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bsr or block_gemm or divide or golden" > gpurun_out/pytest_bsr.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_bsr.log
timeout 600 python tools/time_kernels.py --cases c3d,c4 --maths 1,2 > gpurun_out/time_bsr6.jsonl 2> gpurun_out/time_bsr6.err; echo "rc=$?"; grep bsrmm gpurun_out/time_bsr6.jsonl | cut -c1-330; tail -3 gpurun_out/time_bsr6.err

#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "bsrmm or block_gemm or smoke or golden" > gpurun_out/pytest_bsr.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_bsr.log
timeout 600 python tools/time_kernels.py --cases c3d,c3 --maths 0,1,2 > gpurun_out/time_bsr_tc.log 2>&1; echo "rc=$?"; cut -c1-330 gpurun_out/time_bsr_tc.log

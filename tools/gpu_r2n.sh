#!/bin/bash
# round 2, GPU call N: tile::gather4 CSR kernel -- first contact (small, under a short timeout), then timings on C5
set -x
mkdir -p gpurun_out
timeout 120 python tools/r2_tma_probe.py small > gpurun_out/r2n_tma_small.log 2>&1; echo "rc=$?"
tail -14 gpurun_out/r2n_tma_small.log | cut -c1-300
if grep -q '"equal": true' gpurun_out/r2n_tma_small.log; then
  timeout 600 python tools/r2_tma_probe.py full > gpurun_out/r2n_tma_full.log 2>&1; echo "rc=$?"
  tail -30 gpurun_out/r2n_tma_full.log | cut -c1-200
fi
nvidia-smi --query-gpu=name,memory.used --format=csv

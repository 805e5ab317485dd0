#!/bin/bash
# GPU pass: parity tests, smoke, timing sweep of every kernel on the config shapes.
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
for v in 0 1 2; do
  echo "== csr stream variant $v"
  SPMMDB_CSR_VARIANT=$v timeout 600 python tools/time_kernels.py --cases c2,c2l,c4csr,c5,c5l,c3csr > gpurun_out/time_csr_v$v.log 2>&1
  cat gpurun_out/time_csr_v$v.log | cut -c1-260
done
echo "== rows kernel (stream disabled)"
SPMMDB_CSR_STREAM_MIN_K=100000 timeout 600 python tools/time_kernels.py --cases c2,c4csr,c5l,c3csr > gpurun_out/time_csr_rows.log 2>&1
cat gpurun_out/time_csr_rows.log | cut -c1-260
timeout 900 python tools/time_kernels.py --cases c3,c4 --maths 0 > gpurun_out/time_bsr.log 2>&1; echo "bsr rc=$?"; cut -c1-260 gpurun_out/time_bsr.log

#!/bin/bash
# round 2, GPU call A: new parity tests, vendor comparison, CSR experiments, ncu of the stream kernel on the spec'd C5 graph
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_stream_path.py tests/test_gpu_vendor_parity.py -x -q -m gpu > gpurun_out/r2a_pytest_new.log 2>&1
tail -5 gpurun_out/r2a_pytest_new.log
python tools/r2_csr_experiments.py --out gpurun_out/r2a_csr_experiments.jsonl > gpurun_out/r2a_csr_experiments.log 2>&1
tail -3 gpurun_out/r2a_csr_experiments.log
python baseline/compare.py --out gpurun_out/r2a_vendor.jsonl > gpurun_out/r2a_vendor.log 2>&1
tail -3 gpurun_out/r2a_vendor.log
python bench.py --steps 5 --warmup 3 --no-cpu --e2e-steps 1 > gpurun_out/r2a_bench.log 2>&1
tail -2 gpurun_out/r2a_bench.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:csrmm_stream_kernel -s 6 -c 1 -o gpurun_out/r2a_csr_stream_c5spec \
  python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 --no-extras > gpurun_out/r2a_ncu.log 2>&1
ncu -i gpurun_out/r2a_csr_stream_c5spec.ncu-rep --page raw --csv > gpurun_out/r2a_csr_stream_c5spec_raw.csv 2>/dev/null
rm -f gpurun_out/r2a_csr_stream_c5spec.ncu-rep
python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_stream_path.py --deselect tests/test_gpu_vendor_parity.py > gpurun_out/r2a_pytest_old.log 2>&1
tail -5 gpurun_out/r2a_pytest_old.log

#!/bin/bash
# round 2, GPU call E: lean kernel parity + timing + ncu, BSR column-major epilogue, vendor parity
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_stream_path.py -x -q -m gpu > gpurun_out/r2e_pytest_stream.log 2>&1
tail -5 gpurun_out/r2e_pytest_stream.log
timeout 600 python tools/r2_csr_experiments.py --policies 0 --out gpurun_out/r2e_csr_experiments.jsonl > gpurun_out/r2e_csr_experiments.log 2>&1
tail -3 gpurun_out/r2e_csr_experiments.log
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_drivers.py tests/test_graphio.py -x -q -m gpu -k "histogram or known_answer or microbench or graph" > gpurun_out/r2e_pytest_misc.log 2>&1
tail -3 gpurun_out/r2e_pytest_misc.log
prof() {
  local name=$1 regex=$2; shift 2
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$regex -s 6 -c 1 -o gpurun_out/$name \
    python tools/r2_csr_experiments.py --policies 0 --iters 2 --out gpurun_out/$name.jsonl "$@" > gpurun_out/$name.log 2>&1
  ncu -i gpurun_out/$name.ncu-rep --page raw --csv > gpurun_out/${name}_raw.csv 2>/dev/null
  ncu -i gpurun_out/$name.ncu-rep --page source --csv > gpurun_out/${name}_source.csv 2>/dev/null
  rm -f gpurun_out/$name.ncu-rep
}
prof r2e_bf16_spec_ring1 csrmm_lean_kernel --graphs spec --only "bf16 B bulk 1 rpi 16"
prof r2e_fp32_local_ring1 csrmm_lean_kernel --graphs local --only "fp32 bulk 1 rpi 16"

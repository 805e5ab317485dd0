#!/bin/bash
# round 2, GPU call C: ncu of the CSR kernels where they are NOT DRAM-bound (local graph fp32, bf16 B register / bulk)
set -x
mkdir -p gpurun_out
prof() {  # name, kernel regex, extra env, script args...
  local name=$1 regex=$2; shift 2
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$regex -s 6 -c 1 -o gpurun_out/$name \
    python tools/r2_csr_experiments.py --policies 0 --iters 2 --out gpurun_out/$name.jsonl "$@" > gpurun_out/$name.log 2>&1
  ncu -i gpurun_out/$name.ncu-rep --page raw --csv > gpurun_out/${name}_raw.csv 2>/dev/null
  ncu -i gpurun_out/$name.ncu-rep --page source --csv > gpurun_out/${name}_source.csv 2>/dev/null
  rm -f gpurun_out/$name.ncu-rep
}
prof r2c_fp32_local_stream csrmm_stream_kernel --graphs local --only "fp32 drop-in"
prof r2c_bf16_spec_stream_v1 csrmm_stream_kernel --graphs spec --only "bf16 B variant 1"
prof r2c_bf16_spec_bulk2 csrmm_bulk_kernel --graphs spec --only "bf16 B bulk 2 rpi 16"
prof r2c_fp32_local_bulk1 csrmm_bulk_kernel --graphs local --only "fp32 bulk 1 rpi 16"

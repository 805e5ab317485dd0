#!/bin/bash
# round 2, GPU call J: CUDA-graph replay tests, C2 plain vs graph-launched, ncu launch list + full capture of the bench kernel
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_stream_path.py -x -q -m gpu -k "cuda_graph or two_streams" > gpurun_out/r2j_pytest_graph.log 2>&1
tail -4 gpurun_out/r2j_pytest_graph.log
timeout 600 python tools/r2_c2_graph.py gpurun_out/r2j_c2_graph.jsonl > gpurun_out/r2j_c2_graph.log 2>&1
tail -2 gpurun_out/r2j_c2_graph.log | cut -c1-300
python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 0 --no-extras > gpurun_out/r2j_bench_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2j_launches_bench_c5.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 0 --no-extras > gpurun_out/r2j_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:csrmm_lean_kernel -s 6 -c 1 -o gpurun_out/r2j_csr_lean_c5spec \
  python bench.py --steps 1 --warmup 3 --no-cpu --e2e-steps 0 --no-extras > gpurun_out/r2j_ncu_full.log 2>&1
ncu -i gpurun_out/r2j_csr_lean_c5spec.ncu-rep --page raw --csv > gpurun_out/r2j_csrmm_lean_c5spec_ncu_full_raw.csv 2>/dev/null
rm -f gpurun_out/r2j_csr_lean_c5spec.ncu-rep

#!/bin/bash
# round 2, GPU call O: whole GPU suite on the final tree + default bench + reference arm
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/r2o_pytest_all.log 2>&1
tail -5 gpurun_out/r2o_pytest_all.log
timeout 600 python bench.py > gpurun_out/r2o_bench_default.json 2> gpurun_out/r2o_bench_default.err
tail -c 300 gpurun_out/r2o_bench_default.json
timeout 300 python bench.py --workload c3_proteins_gorder --op hybrid32 --density 0.03 --cpu-budget 5 > gpurun_out/r2o_bench_c3_hybrid32.json 2>&1
tail -c 300 gpurun_out/r2o_bench_c3_hybrid32.json

#!/bin/bash
# round 2, GPU call F: whole GPU suite, tf32 peak, config-3 pipeline, bench lines for the configs, vendor C4 block GEMM
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2f_pytest_all.log 2>&1
tail -5 gpurun_out/r2f_pytest_all.log
python tools/measure_tf32_peak.py gpurun_out/r2_tf32_peak.json > gpurun_out/r2f_tf32.log 2>&1
tail -1 gpurun_out/r2f_tf32.log
timeout 1200 python tools/r2_c3_pipeline.py --out gpurun_out/r2f_c3_pipeline.jsonl > gpurun_out/r2f_c3_pipeline.log 2>&1
tail -3 gpurun_out/r2f_c3_pipeline.log | cut -c1-300
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2f_bench_default.json 2> gpurun_out/r2f_bench_default.err
tail -c 600 gpurun_out/r2f_bench_default.json
timeout 600 python baseline/compare.py --configs c4 --out gpurun_out/r2f_vendor_c4.jsonl > gpurun_out/r2f_vendor_c4.log 2>&1
for w in c1_uniform:csr c2_arxiv:csr c4_molhiv:csr c4_molhiv:bsr16 c4_molhiv:bsr32 c4_molhiv:block_gemm32 c5_products:csr_bf16b; do
  timeout 600 python bench.py --workload ${w%%:*} --op ${w##*:} --steps 10 --warmup 3 --cpu-budget 5 >> gpurun_out/r2f_bench_configs.jsonl 2>> gpurun_out/r2f_bench_configs.err
done
wc -l gpurun_out/r2f_bench_configs.jsonl

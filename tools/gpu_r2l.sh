#!/bin/bash
# round 2, GPU call L: whole GPU suite, config-3 bench lines through bench.py (Gorder / RCM orderings, hybrid object), default bench
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/r2l_pytest_all.log 2>&1
tail -5 gpurun_out/r2l_pytest_all.log
rm -f gpurun_out/r2l_bench_c3.jsonl
for w in c3_proteins_gorder:hybrid32 c3_proteins_gorder:hybrid16 c3_proteins_gorder:csr c3_proteins_gorder:bsr32 c3_proteins_rcm:hybrid32 c3_proteins_shuffled:csr; do
  timeout 900 python bench.py --workload ${w%%:*} --op ${w##*:} --density 0.03 --steps 10 --warmup 3 --cpu-budget 5 >> gpurun_out/r2l_bench_c3.jsonl 2>> gpurun_out/r2l_bench_c3.err
done
wc -l gpurun_out/r2l_bench_c3.jsonl; tail -3 gpurun_out/r2l_bench_c3.err
timeout 600 python bench.py > gpurun_out/r2l_bench_default.json 2> gpurun_out/r2l_bench_default.err
tail -c 400 gpurun_out/r2l_bench_default.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2l_bench_reference.json 2>&1
tail -c 300 gpurun_out/r2l_bench_reference.json
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2l_smoke.log 2>&1; tail -2 gpurun_out/r2l_smoke.log

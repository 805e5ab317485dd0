#!/bin/bash
# round-2 run V: config 3 reordered ON THE DEVICE (spmmdb_order_rcm) next to the reference's CPU RCM order
mkdir -p gpurun_out
rm -f gpurun_out/r2v_bench_c3_rcm_gpu.jsonl
for w in c3_proteins_rcm_gpu:hybrid32 c3_proteins_rcm_gpu:hybrid16 c3_proteins_rcm_gpu:csr c3_proteins_rcm_gpu:bsr32 c3_proteins_rcm:hybrid32 c3_proteins_rcm:csr; do
  timeout 600 python bench.py --workload ${w%%:*} --op ${w##*:} --density 0.03 --steps 10 --warmup 3 --no-cpu >> gpurun_out/r2v_bench_c3_rcm_gpu.jsonl 2>> gpurun_out/r2v.err
done
python - <<P
import json
for l in open("gpurun_out/r2v_bench_c3_rcm_gpu.jsonl"):
    d=json.loads(l)
    print(d["config"]["workload"][:40], d["config"]["op"], round(d["ms_per_step"],3), {k:d[k] for k in d if k.endswith("_once") or k.startswith("unpermuted") or k in ("block_part","hybrid")})
P
tail -3 gpurun_out/r2v.err

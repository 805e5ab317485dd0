#!/bin/bash
# round-2 run Q: permutation application after the register-sort rewrite (tests, timings, launch list), and the
# host-buffer path's slab plan (HOST_TAIL) on the products-shaped graph
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k permute > gpurun_out/perm_tests.log 2>&1; tail -3 gpurun_out/perm_tests.log
python tools/time_kernels.py --cases perm > gpurun_out/r2q_permute.jsonl 2>gpurun_out/perm_err.log; cut -c1-200 gpurun_out/r2q_permute.jsonl
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:perm_ --csv --log-file gpurun_out/r2q_permute_launches.csv \
  python tools/time_kernels.py --cases perm > /dev/null 2>&1
: > gpurun_out/r2q_host_slabs.jsonl
for cfg in "64 0" "32 0" "16 0" "16 8" "0 8"; do
  set -- $cfg
  SPMMDB_HOST_TAIL=$1 SPMMDB_HOST_SLABS=$2 python bench.py --steps 3 --warmup 3 --e2e-steps 5 --no-cpu --no-extras 2>gpurun_out/host_err.log |
    python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print(json.dumps(dict(host_tail=$1, host_slabs=$2, e2e=d['e2e'], ms_per_step=d['ms_per_step'])))" | tee -a gpurun_out/r2q_host_slabs.jsonl
done
python -m pytest tests -x -q -m gpu -k "host" 2>&1 | tail -3

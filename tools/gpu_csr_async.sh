#!/bin/bash
mkdir -p gpurun_out
for v in 1 2 3; do
SPMMDB_CSR_ASYNC=$v timeout 600 python -m pytest tests -m gpu -x -q -k "csrmm or check_result or shard or smoke" > gpurun_out/pytest_async$v.log 2>&1; echo "async=$v pytest rc=$?"; tail -3 gpurun_out/pytest_async$v.log
done
for v in 0 1 2 3; do
  echo "== async $v"
  SPMMDB_CSR_ASYNC=$v timeout 600 python tools/time_kernels.py --cases c2,c2l,c4csr,c5,c5l 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l.rstrip()); continue
    print(d['case'], d['K'], 'ms_med %.4f ms_min %.4f gflops %.0f frac %.3f' % (d['ms_med'], d['ms_min'], d['gflops'], d['frac']))
"
done

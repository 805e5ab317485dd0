#!/bin/bash
mkdir -p gpurun_out
for v in 0 1; do for r in 1 2 4 8 16 32; do
  echo "== variant $v rpi $r"
  SPMMDB_CSR_VARIANT=$v SPMMDB_CSR_RPI=$r timeout 600 python tools/time_kernels.py --cases c2,c5l,c5,c4csr 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l.rstrip()); continue
    print(d['case'], d['K'], 'ms_med %.4f ms_min %.4f gflops %.0f frac %.3f' % (d['ms_med'], d['ms_min'], d['gflops'], d['frac']))
"
done; done

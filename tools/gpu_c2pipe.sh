#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_shard.py -m gpu -q -x -k "csrmm or csr or shard or gespmm or hybrid or divide" 2>&1 | tail -3
for cfg in "0 3 4 64" "1 3 4 64" "1 3 4 32" "1 6 4 64" "2 3 4 64"; do
  set -- $cfg
  echo "== PIPE=$1 PIPE_CTAS=$2 LONG_CTAS=$3 SEG=$4"
  SPMMDB_CSR_PIPE=$1 SPMMDB_CSR_PIPE_CTAS=$2 SPMMDB_CSR_LONG_CTAS=$3 SPMMDB_CSR_SEG=$4 timeout 300 python tools/time_kernels.py --cases c2,c2l 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['case'],'ms_med %.4f ms_min %.4f'%(d['ms_med'],d['ms_min']))"
done

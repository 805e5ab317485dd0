#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bsr" > gpurun_out/pytest_bsr.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_bsr.log
timeout 600 python tools/time_kernels.py --cases c3d,c4 --maths 1,3 2>/dev/null | grep bsrmm | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['case'],'bs',d['bs'],'math',d['math'],'ms %.3f dense %.1f TF useful %.0f GF frac_hbm %.3f'%(d['ms_med'],d['dense_tflops'],d['useful_gflops'],d['frac_hbm']))"

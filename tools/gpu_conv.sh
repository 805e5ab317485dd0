#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "csr2bsr or bsr2csr or divide or permute or conversion or round" > gpurun_out/pytest_conv.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_conv.log
timeout 600 python tools/time_kernels.py --cases c3d,c4,perm --maths 1 2>/dev/null | grep -v '"bsrmm"\|"csrmm"' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if d['kernel']=='permute_csr': print(d['case'],'permute ms %.3f'%d['ms_med'])
    else: print(d['case'],d['kernel'],'bs',d['bs'],'ms %.3f frac %.3f'%(d['ms_med'],d['frac']))"

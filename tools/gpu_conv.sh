#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/time_kernels.py --cases c3d,c3,c4,c3h --maths 1 > gpurun_out/r1m_conv.jsonl 2>/dev/null; grep -v '"bsrmm"\|"csrmm"\|bsr2csr' gpurun_out/r1m_conv.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if d['kernel']=='hybrid': print(d['case'],'divide bs',d['bs'],'dens',d['density'],'ms %.3f hybrid %.3f csr %.3f'%(d['ms_divide'],d['ms_hybrid'],d['ms_csr_only']))
    else: print(d['case'],d['kernel'],'bs',d['bs'],'ms %.3f frac %.3f'%(d['ms_med'],d['frac']))"

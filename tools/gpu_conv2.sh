#!/bin/bash
timeout 900 python -m pytest tests -m gpu -q -x -k "csr2bsr or bsr2csr or divide or conversion or metrics or check_result or run_bsrmm or hybrid" 2>&1 | tail -3
timeout 600 python tools/time_kernels.py --cases c3d,c3,c4,c3h --maths 1 2>/dev/null | tee gpurun_out/r1r_conv.jsonl | grep -v '"bsrmm"\|"csrmm"\|bsr2csr' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if d['kernel']=='hybrid': print(d['case'],'divide bs',d['bs'],'dens',d['density'],'ms %.3f'%(d['ms_divide']))
    else: print(d['case'],d['kernel'],'bs',d['bs'],'ms %.3f frac %.3f'%(d['ms_med'],d['frac']))"

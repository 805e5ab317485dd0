#!/bin/bash
for c in 0 6 4 3 2; do
echo "== BR_CTAS=$c"
SPMMDB_CONVERT_BR_CTAS=$c timeout 600 python tools/time_kernels.py --cases c3d,c3h --maths 1 2>/dev/null | grep -v '"bsrmm"\|"csrmm"\|bsr2csr' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if d['kernel']=='hybrid':
        if d['density'] in (0.05,): print(d['case'],'divide bs',d['bs'],'dens',d['density'],'ms %.3f'%(d['ms_divide']))
    else: print(d['case'],d['kernel'],'bs',d['bs'],'ms %.3f frac %.3f'%(d['ms_med'],d['frac']))"
done

#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "bsr2csr or roundtrip or check_result" 2>&1 | tail -4
for sc in 1 0; do
echo "== SPMMDB_BSR2CSR_SCALAR=$sc"
SPMMDB_BSR2CSR_SCALAR=$sc timeout 600 python tools/time_kernels.py --cases c3d,c3,c4 --maths 1 2>/dev/null | tee gpurun_out/r1n_bsr2csr_scalar$sc.jsonl | grep bsr2csr | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['case'],d['kernel'],'bs',d['bs'],'ms %.3f gbs %.0f frac %.3f'%(d['ms_med'],d['gbs'],d['frac']))"
done

#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
timeout 900 python tools/time_kernels.py --cases c3h --maths 1 > gpurun_out/time_hybrid.jsonl 2> gpurun_out/time_hybrid.err; echo "rc=$?"; python - <<'PY'
import json
for l in open('gpurun_out/time_hybrid.jsonl'):
    d=json.loads(l); print(d['case'],d['bs'],d['density'],'dense_part %.2f fill %.3f hybrid %.3f csr %.3f speedup %.2f divide %.2f err %.1e'%(d['dense_part'],d['block_fill'],d['ms_hybrid'],d['ms_csr_only'],d['speedup'],d['ms_divide'],d['rel_diff_vs_csr']))
PY

#!/bin/bash
mkdir -p gpurun_out
for s in 1 2 4 8; do echo "slabs=$s"; SPMMDB_HOST_SLABS=$s timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --e2e-steps 3 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e'])"; done

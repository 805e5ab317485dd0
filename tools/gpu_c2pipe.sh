#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_shard.py -m gpu -q -x -k "csrmm or csr or shard or gespmm or hybrid or divide" 2>&1 | tail -3
for s in 1 0; do echo "== SMALL=$s"; SPMMDB_CSR_SMALL=$s python tools/c2_probe.py 2>&1 | grep "preferential\|uniform"; done

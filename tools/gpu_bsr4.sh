#!/bin/bash
for d in 32 39; do for P in 1 3; do echo "== skip=$d P=$P"; SPMMDB_BSR_PRODUCERS=$P SPMMDB_BSR_DEBUG_SKIP=$d timeout 100 python tools/bsr_one.py --bs 32 --math 1 --iters 0 2>&1 | tail -12; done; done

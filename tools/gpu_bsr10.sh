#!/bin/bash
for bs in 32 16; do for cfg in "2 128 3" "1 256 3" "1 128 3" "2 128 2" "2 128 4" "2 64 3"; do set -- $cfg
echo "bs=$bs cps=$1 NT=$2 P=$3 $(SPMMDB_BSR_CTAS_PER_SM=$1 SPMMDB_BSR_NT=$2 SPMMDB_BSR_PRODUCERS=$3 timeout 120 python tools/bsr_one.py --bs $bs --math 1 --iters 7 2>/dev/null | cut -c1-160)"
done; done

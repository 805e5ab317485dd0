#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/tc_debug.py > gpurun_out/tc_debug.log 2>&1; echo "debug rc=$? ok=$(grep -c 'bad 0/' gpurun_out/tc_debug.log)"; grep -v "bad 0/" gpurun_out/tc_debug.log | head -5
SPMMDB_BSR_CTAS_PER_SM=1 timeout 120 python tools/tc_debug.py > gpurun_out/tc_debug1.log 2>&1; echo "debug(cps1) rc=$? ok=$(grep -c 'bad 0/' gpurun_out/tc_debug1.log)"; grep -v "bad 0/" gpurun_out/tc_debug1.log | head -5
for bs in 32 16; do for cfg in "1 128" "1 256" "2 128" "2 64"; do set -- $cfg
SPMMDB_BSR_CTAS_PER_SM=$1 SPMMDB_BSR_NT=$2 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"bsrmm_tc" -s 1 -c 1 --csv --log-file gpurun_out/l5.csv python tools/bsr_one.py --bs $bs --math 1 --iters 1 > /dev/null 2>&1
echo "bs=$bs cps=$1 NT=$2 kernel_ns=$(tail -1 gpurun_out/l5.csv | awk -F'","' '{print $NF}')"
done; done

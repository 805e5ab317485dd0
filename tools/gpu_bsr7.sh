#!/bin/bash
# where does the tcgen05 BSR kernel spend its time after the elect.sync change (kernel-only ns via ncu, plus clock64 counters)
mkdir -p gpurun_out
for bs in 32 16; do for math in 1 2; do for dbg in 0 3 4 7 8; do
SPMMDB_BSR_DEBUG_SKIP=$dbg ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"bsrmm_tc_kernel" -s 1 -c 1 --csv --log-file gpurun_out/l7.csv python tools/bsr_one.py --bs $bs --math $math --iters 1 > /dev/null 2>&1
echo "bs=$bs math=$math dbg=$dbg kernel_ns=$(tail -1 gpurun_out/l7.csv | awk -F'","' '{print $NF}')"
done; done; done
for bs in 32 16; do
SPMMDB_BSR_DEBUG_SKIP=32 timeout 120 python tools/bsr_one.py --bs $bs --math 1 --iters 1 2>&1 | grep -E "mma:|producer" | sort | uniq -c | sort -rn | head -6
done

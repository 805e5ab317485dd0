#!/bin/bash
mkdir -p gpurun_out
for bs in 32; do for math in 1 2; do for dbg in 3 67 0 64; do
SPMMDB_BSR_DEBUG_SKIP=$dbg ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"bsrmm_tc_kernel" -s 1 -c 1 --csv --log-file gpurun_out/l8.csv python tools/bsr_one.py --bs $bs --math $math --iters 1 > /dev/null 2>&1
echo "bs=$bs math=$math dbg=$dbg kernel_ns=$(tail -1 gpurun_out/l8.csv | awk -F'","' '{print $NF}')"
done; done; done
SPMMDB_BSR_DEBUG_SKIP=32 timeout 120 python tools/bsr_one.py --bs 32 --math 1 --iters 1 2>&1 | grep -E "mma:" | head -4

#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/bsr_one.py --bs 32 --math 1 --iters 2"
$CMD > gpurun_out/bsr_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:bsrmm_tc -s 2 -c 1 -f -o gpurun_out/bsr_tc_r1 $CMD > gpurun_out/ncu_bsr.log 2>&1
echo "rc=$?"; cat gpurun_out/bsr_plain.log; tail -3 gpurun_out/ncu_bsr.log

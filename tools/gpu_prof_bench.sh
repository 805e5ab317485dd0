#!/bin/bash
# ncu evidence for the bench command: launch list (per-launch device time) + one full capture of the dominant kernel
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 1"
$CMD > gpurun_out/bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"csrmm" -c 40 --csv --log-file gpurun_out/r1d_launches_bench.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "launch rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"csrmm_stream" -s 3 -c 1 -f -o gpurun_out/r1d_csrmm_stream_c5 $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/ncu_full.log

#!/bin/bash
# ncu evidence for the bench command: launch list (per-launch device time) + one full capture of the dominant kernel
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --e2e-steps 1"
$CMD > gpurun_out/bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"csrmm" -c 40 --csv --log-file gpurun_out/r1i_launches_bench.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "launch rc=$?"
# launches per step: stream kernel + its sweep launch; skip the 3 warm-up steps
ncu --set full --clock-control none --import-source on -k regex:"csrmm_stream" -s 6 -c 1 -f -o gpurun_out/r1i_csrmm_stream_c5 $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/ncu_full.log
ncu -i gpurun_out/r1i_csrmm_stream_c5.ncu-rep --page raw --csv > gpurun_out/r1i_csrmm_stream_c5_raw.csv 2>/dev/null; wc -c gpurun_out/r1i_csrmm_stream_c5_raw.csv
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-1200 gpurun_out/bench.log

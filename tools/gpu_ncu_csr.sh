#!/bin/bash
# ncu pass for the CSR kernel on the bench workload (c5 local) + c4 + c2.
mkdir -p gpurun_out
CMD="python tools/time_kernels.py --cases c5l,c4csr,c2"
$CMD > gpurun_out/ncu_plain.log 2>&1 || { echo plain run failed; tail gpurun_out/ncu_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:csrmm_rows -c 6 -o gpurun_out/csrmm_r1 -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"; tail -3 gpurun_out/ncu_launch.log

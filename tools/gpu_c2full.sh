#!/bin/bash
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:rows_pipe -s 6 -c 1 -o gpurun_out/c2_rows_pipe -f python tools/time_kernels.py --cases c2 > gpurun_out/c2_full.log 2>&1
echo rc=$?
ncu -i gpurun_out/c2_rows_pipe.ncu-rep --page raw --csv > gpurun_out/c2_rows_pipe_raw.csv 2>/dev/null
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/c2_rows_pipe_raw.csv')))
h,u,v=rows[0],rows[1],rows[2]
keys=['gpu__time_duration.sum','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct','dram__throughput.avg.pct_of_peak_sustained_elapsed','lts__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','smsp__warps_eligible.avg.per_cycle_active','launch__registers_per_thread','launch__occupancy_limit_registers','smsp__cycles_active.avg','sm__cycles_elapsed.max']
for i,k in enumerate(h):
    if k in keys or 'smsp__average_warp' in k and 'issue_stalled' in k and 'not_issued' not in k and float(v[i].replace(',','') or 0)>0.3:
        print(k,u[i],v[i])
PY

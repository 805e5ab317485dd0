#!/bin/bash
# ncu --set full of the current conversion kernels on the dense-community C3 graph (one launch each)
mkdir -p gpurun_out
CMD="python tools/time_kernels.py --cases c3d --maths 1"
for k in bsr2csr_vec csr2bsr_br; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 3 -c 2 -o gpurun_out/r1s_$k -f $CMD > gpurun_out/r1s_$k.log 2>&1
  echo "$k rc=$?"
  ncu -i gpurun_out/r1s_$k.ncu-rep --page raw --csv > gpurun_out/r1s_${k}_raw.csv 2>/dev/null
done
python - <<'PY'
import csv
keys=['gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','dram__throughput.avg.pct_of_peak_sustained_elapsed','lts__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio','smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio','smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio','smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio','smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio','lts__t_sector_hit_rate.pct']
for k in ('bsr2csr_vec','csr2bsr_br'):
    rows=list(csv.reader(open(f'gpurun_out/r1s_{k}_raw.csv')))
    h,u=rows[0],rows[1]
    for r in rows[2:]:
        name=r[h.index('Kernel Name')][:60]
        print('##',name,r[h.index('Grid Size')] if 'Grid Size' in h else '')
        for kk in keys:
            if kk in h: print('  ',kk,u[h.index(kk)],r[h.index(kk)])
PY

#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/time_kernels.py --cases c2"
$CMD > gpurun_out/c2_plain.log 2>&1; cat gpurun_out/c2_plain.log | cut -c1-200
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:csrmm -c 18 --csv --log-file gpurun_out/c2_launches.csv $CMD > gpurun_out/c2_ncu.log 2>&1
echo "rc=$?"
python - <<'PY'
import csv
rows=list(csv.reader(l for l in open('gpurun_out/c2_launches.csv') if l.startswith('"')))
h=rows[0]; i={k:h.index(k) for k in ('ID','Kernel Name','Metric Name','Metric Value','Grid Size')}
out={}
for r in rows[1:]:
    out.setdefault((r[i['ID']],r[i['Kernel Name']][:60],r[i['Grid Size']]),{})[r[i['Metric Name']]]=r[i['Metric Value']]
for k,v in list(out.items())[-6:]:
    print(k,v)
PY

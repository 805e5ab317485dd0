#!/bin/bash
# launch list (device time + DRAM bytes) of the conversion kernels on the dense-community C3 graph
mkdir -p gpurun_out
CMD="python tools/time_kernels.py --cases c3d,c3h --maths 1"
$CMD > gpurun_out/conv_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"bsr2csr|csr2bsr|divide" --csv --log-file gpurun_out/conv_launches.csv $CMD > gpurun_out/conv_ncu.log 2>&1
echo rc=$?
python - <<'PY'
import csv, json, collections
rows=list(csv.reader(l for l in open('gpurun_out/conv_launches.csv') if l.startswith('"')))
h=rows[0]; i={k:h.index(k) for k in ('ID','Kernel Name','Metric Name','Metric Value','Grid Size')}
out=collections.OrderedDict()
for r in rows[1:]:
    out.setdefault((r[i['ID']],r[i['Kernel Name']],r[i['Grid Size']]),{})[r[i['Metric Name']]]=float(r[i['Metric Value']].replace(',',''))
agg=collections.OrderedDict()
for (id_,k,g),v in out.items():
    key=(k[:80],g)
    a=agg.setdefault(key,dict(n=0,us=[],rd=0,wr=0,wa=0))
    a['n']+=1; a['us'].append(v['gpu__time_duration.sum']/1e3); a['rd']=v['dram__bytes_read.sum']/1e6; a['wr']=v['dram__bytes_write.sum']/1e6; a['wa']=v['sm__warps_active.avg.pct_of_peak_sustained_active']
with open('gpurun_out/r1p_conversion_launches.jsonl','w') as f:
    for (k,g),a in agg.items():
        us=sorted(a['us'])[len(a['us'])//2]
        d=dict(kernel=k,grid=g,launches=a['n'],us_med=round(us,2),dram_read_MB=round(a['rd'],1),dram_write_MB=round(a['wr'],1),dram_GBs=round((a['rd']+a['wr'])/us*1e3/1e3,1),warps_active_pct=round(a['wa'],1))
        f.write(json.dumps(d)+"\n"); print(d)
PY

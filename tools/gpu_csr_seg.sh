#!/bin/bash
# segment-length sweep + DRAM traffic of the stream kernel with inline segments
mkdir -p gpurun_out
: > gpurun_out/seg_sweep.txt
for seg in 64 128 256 512; do echo "seg=$seg" >> gpurun_out/seg_sweep.txt; SPMMDB_CSR_SEG=$seg timeout 300 python tools/time_kernels.py --cases c2,c2l >> gpurun_out/seg_sweep.txt 2>/dev/null; done
for seg in 256 512 1024 2048; do echo "seg=$seg" >> gpurun_out/seg_sweep.txt; SPMMDB_CSR_SEG=$seg timeout 300 python tools/time_kernels.py --cases c5l >> gpurun_out/seg_sweep.txt 2>/dev/null; done
cut -c1-200 gpurun_out/seg_sweep.txt
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:"csrmm" -c 8 --csv --log-file gpurun_out/r1f_c5l_traffic.csv python tools/time_kernels.py --cases c5l > gpurun_out/ncu_c5l.log 2>&1
echo "ncu rc=$?"; grep -v "^==" gpurun_out/r1f_c5l_traffic.csv | cut -d, -f5,13,15 | tail -42

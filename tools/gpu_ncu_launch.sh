#!/bin/bash
# launch list (per-kernel device time) of the CSR path on c2 / c5l / c3
mkdir -p gpurun_out
CMD="python tools/time_kernels.py --cases c2,c5l,c3csr"
SPMMDB_CSR_VARIANT=1 $CMD > gpurun_out/ncu_plain.log 2>&1 &&
SPMMDB_CSR_VARIANT=1 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:csrmm --csv --log-file gpurun_out/launches_csr.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_launch.log

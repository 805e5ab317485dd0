#!/bin/bash
# ncu --set full of the tcgen05 BSR kernel (C3-dense, tf32, bs 32 and 16) and of the conversion kernels
mkdir -p gpurun_out
for bs in 32 16; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"bsrmm_tc_kernel" -s 1 -c 1 -f -o gpurun_out/r1l_bsrmm_tc_bs$bs python tools/bsr_one.py --bs $bs --math 1 --iters 1 > gpurun_out/ncu_bsr_$bs.log 2>&1; echo "bs=$bs rc=$?"
ncu -i gpurun_out/r1l_bsrmm_tc_bs$bs.ncu-rep --page raw --csv > gpurun_out/r1l_bsrmm_tc_bs${bs}_raw.csv 2>/dev/null
done
timeout 600 ncu --set full --clock-control none -k regex:"csr2bsr_kernel|bsr2csr_kernel|bsr_plan_kernel" -c 6 -f -o gpurun_out/r1l_convert python tools/time_kernels.py --cases c3d --maths 1 > gpurun_out/ncu_conv.log 2>&1; echo "conv rc=$?"
ncu -i gpurun_out/r1l_convert.ncu-rep --page raw --csv > gpurun_out/r1l_convert_raw.csv 2>/dev/null
ls -la gpurun_out/r1l_*

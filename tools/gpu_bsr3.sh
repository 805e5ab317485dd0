#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/bsr_one.py --bs 32 --iters 1"
$CMD --math 2 > gpurun_out/plain3.log 2>&1 && ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg,sm__cycles_elapsed.max,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,smsp__inst_executed.sum --clock-control none -k regex:"bsrmm_tc|bsr_plan|cvt_bf16|prep_b" -s 4 -c 4 --csv --log-file gpurun_out/l3_bf16.csv $CMD --math 2 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg,sm__cycles_elapsed.max,smsp__inst_executed.sum --clock-control none -k regex:"bsrmm_tc|bsr_plan|cvt_bf16|prep_b" -s 2 -c 2 --csv --log-file gpurun_out/l3_tf32.csv $CMD --math 1 > /dev/null 2>&1
SPMMDB_BSR_NT=256 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg,sm__cycles_elapsed.max,smsp__inst_executed.sum --clock-control none -k regex:"bsrmm_tc|bsr_plan|cvt_bf16|prep_b" -s 2 -c 2 --csv --log-file gpurun_out/l3_tf32_nt256.csv $CMD --math 1 > /dev/null 2>&1
python tools/bsr_one.py --bs 16 --iters 1 --math 2 > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg,sm__cycles_elapsed.max,smsp__inst_executed.sum --clock-control none -k regex:"bsrmm_tc|bsr_plan|cvt_bf16|prep_b" -s 4 -c 4 --csv --log-file gpurun_out/l3_bf16_16.csv python tools/bsr_one.py --bs 16 --iters 1 --math 2 > /dev/null 2>&1
for f in gpurun_out/l3_*.csv; do echo == $f; grep -v "^==" $f | python -c "
import sys,csv
r=list(csv.reader(sys.stdin)); h=r[0]
for x in r[1:]:
    d=dict(zip(h,x)); print(d['Kernel Name'][:60].split('(')[0][-40:], d['Metric Name'], d['Metric Value'])
"; done

#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/tc_debug.py > gpurun_out/tc_debug.log 2>&1; echo "debug rc=$?"; grep -c "bad 0/" gpurun_out/tc_debug.log; grep -v "bad 0/" gpurun_out/tc_debug.log | head
for P in 1 2 3 4; do for bs in 16 32; do
SPMMDB_BSR_PRODUCERS=$P timeout 200 python tools/bsr_one.py --bs $bs --math 1
done; done
SPMMDB_BSR_B3D=0 timeout 200 python tools/bsr_one.py --bs 32 --math 1
SPMMDB_BSR_NT=256 timeout 200 python tools/bsr_one.py --bs 32 --math 1
SPMMDB_BSR_NT=256 timeout 200 python tools/bsr_one.py --bs 16 --math 1
timeout 200 python tools/bsr_one.py --bs 32 --math 2
timeout 200 python tools/bsr_one.py --bs 64 --math 2

#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/time_kernels.py --cases c3h,c3,c4 --maths 0,1 > gpurun_out/time_hybrid.jsonl 2> gpurun_out/time_hybrid.err; echo "rc=$?"; cut -c1-420 gpurun_out/time_hybrid.jsonl; tail -5 gpurun_out/time_hybrid.err

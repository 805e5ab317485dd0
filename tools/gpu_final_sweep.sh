#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python tools/time_kernels.py --cases c2,c2l,c4,c5,c5l,c3,c3d,perm --maths 0,1 > gpurun_out/r1j_sweep.jsonl 2> gpurun_out/r1j_sweep.err; echo "rc=$?"; cut -c1-260 gpurun_out/r1j_sweep.jsonl; tail -3 gpurun_out/r1j_sweep.err

#!/bin/bash
# round-2 run U: validation of the head -- full GPU suite, smoke, the default bench line and the reference arm
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r2u_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2u_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
python bench.py > gpurun_out/r2u_bench_default.json 2>gpurun_out/r2u_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2u_bench_reference.json 2>>gpurun_out/r2u_bench.err; echo "ref rc=$?"
python - <<P
import json
d=json.loads(open("gpurun_out/r2u_bench_default.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("value","ms_per_step","gpu_launches","clocks")}, d["roofline"]["frac"], d["e2e"]["ms_per_step"], d["e2e"]["matches_device_path"], d.get("cpu_baseline"))
r=json.loads(open("gpurun_out/r2u_bench_reference.json").read().strip().splitlines()[-1])
print(r["impl"], r["value"], r["ms_per_step"], r["cpu_baseline"])
P

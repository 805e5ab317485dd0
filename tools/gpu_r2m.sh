#!/bin/bash
# round 2, GPU call M: ncu --set full of the tcgen05 BSR kernel on C4 (bs 32), the Gorder-reordered C3 graph (plain BSR
# bs 32 and the hybrid split's block part bs 32 / 16) -> profiles/r2_bsr_ncu.md
set -x
mkdir -p gpurun_out
prof() {
  local name=$1; shift
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:bsrmm_tc_kernel -s 4 -c 1 -o gpurun_out/$name \
    python bench.py "$@" --steps 2 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/$name.log 2>&1
  ncu -i gpurun_out/$name.ncu-rep --page raw --csv > gpurun_out/${name}_raw.csv 2>/dev/null
  rm -f gpurun_out/$name.ncu-rep
}
prof r2m_bsr_c4_bs32 --workload c4_molhiv --op bsr32
prof r2m_bsr_c3gorder_hybrid32 --workload c3_proteins_gorder --op hybrid32 --density 0.03
prof r2m_bsr_c3gorder_hybrid16 --workload c3_proteins_gorder --op hybrid16 --density 0.03
prof r2m_bsr_c3gorder_bsr32 --workload c3_proteins_gorder --op bsr32
ls -la gpurun_out | grep r2m

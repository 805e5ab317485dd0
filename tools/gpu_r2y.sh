#!/bin/bash
# round-2 run Y: ncu --set full of the permutation sort kernels (C3: the 16-keys-per-lane warp kernel; C5: the CTA
# kernel on the 513..4096-entry rows) and of the BFS expand kernels on C5
mkdir -p gpurun_out
cat > /tmp/perm_c3.py <<'P'
import sys, torch
sys.path.insert(0, ".")
import spmm_denseblock_b200 as pkg
from spmm_denseblock_b200 import graphs
cfg = sys.argv[1]
rp, ci = graphs.make_config_graph(cfg, device="cuda")
n = rp.numel() - 1
perm = torch.randperm(n, device="cuda").to(torch.int32)
val = torch.ones(ci.numel(), device="cuda")
for _ in range(2):
    pkg.permute_csr(rp, ci, val, perm)
if len(sys.argv) > 2:
    for _ in range(2):
        pkg.order_bfs(rp, ci)
torch.cuda.synchronize()
P
ncu --set full --clock-control none -k regex:perm_warp_kernel -s 1 -c 1 -o gpurun_out/r2y_perm_warp16_c3 python /tmp/perm_c3.py c3_proteins > /dev/null 2>&1
ncu --set full --clock-control none -k regex:perm_cta_kernel -s 3 -c 2 -o gpurun_out/r2y_perm_cta_c5 python /tmp/perm_c3.py c5_products > /dev/null 2>&1
ncu --set full --clock-control none -k regex:perm_warp1_kernel -s 1 -c 1 -o gpurun_out/r2y_perm_warp1_c5 python /tmp/perm_c3.py c5_products > /dev/null 2>&1
for f in r2y_perm_warp16_c3 r2y_perm_cta_c5 r2y_perm_warp1_c5; do
  ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/${f}_ncu_full_raw.csv 2>/dev/null
  rm -f gpurun_out/$f.ncu-rep
done
ls -la gpurun_out/r2y_*

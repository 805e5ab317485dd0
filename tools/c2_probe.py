"""C2 probes: is the short-row regime bound by bytes, by hub-column hot spots, or by per-row latency?"""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import spmm_denseblock_b200 as pkg
from spmm_denseblock_b200 import graphs
from tools.time_kernels import timeit

flush = torch.empty(256 << 20, dtype=torch.float32, device="cuda")
rp, ci = graphs.make_config_graph("c2_arxiv", device="cuda")
n, nnz = rp.numel() - 1, ci.numel()
deg = (rp[1:] - rp[:-1]).long()
cases = {"preferential": ci}
g = torch.Generator(device="cuda"); g.manual_seed(7)
cases["uniform_cols"] = torch.randint(0, n, (nnz,), device="cuda", generator=g, dtype=torch.int32)
cases["same_col_per_row"] = torch.arange(n, device="cuda", dtype=torch.int32).repeat_interleave(deg)
short = deg <= 64
for name, cols in cases.items():
    for K in (64, 128, 256):
        val = torch.ones(nnz, device="cuda")
        B = torch.rand(n, K, device="cuda") * 2 - 1
        C = torch.empty(n, K, device="cuda")
        med, best = timeit(lambda: pkg.csrmm(rp, cols, val, B, C, n, n, K, nnz=nnz), flush=flush)
        print(json.dumps(dict(case=name, K=K, ms_med=round(med, 4), ms_min=round(best, 4))), flush=True)
# only short rows (hubs truncated to 64 entries): the rows kernel alone
d2 = torch.clamp(deg, max=64)
rp2 = torch.zeros(n + 1, dtype=torch.int32, device="cuda"); rp2[1:] = torch.cumsum(d2, 0).int()
keep = (torch.arange(nnz, device="cuda") - rp[:-1].long().repeat_interleave(deg)) < 64
ci2 = ci[keep].contiguous()
for K in (128,):
    val = torch.ones(ci2.numel(), device="cuda"); B = torch.rand(n, K, device="cuda"); C = torch.empty(n, K, device="cuda")
    med, best = timeit(lambda: pkg.csrmm(rp2, ci2, val, B, C, n, n, K, nnz=ci2.numel()), flush=flush)
    print(json.dumps(dict(case="hubs_truncated_to_64", nnz=ci2.numel(), K=K, ms_med=round(med, 4), ms_min=round(best, 4))), flush=True)

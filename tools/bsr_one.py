"""One BSR tensor-core case for profiling: c3 block-dense graph, given bs / math."""
import argparse, os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import spmm_denseblock_b200 as pkg
from spmm_denseblock_b200 import graphs
ap = argparse.ArgumentParser()
ap.add_argument("--bs", type=int, default=32)
ap.add_argument("--math", type=int, default=1)
ap.add_argument("--K", type=int, default=128)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--p_in", type=float, default=1.0)
ap.add_argument("--scale", type=float, default=1.0)
a = ap.parse_args()
rp, ci = graphs.make_config_graph("c3_proteins", device="cuda", scale=a.scale, p_in=a.p_in)
n = rp.numel() - 1
val = torch.ones(ci.numel(), device="cuda")
bs, K = a.bs, a.K
brp, bci, bv = pkg.csr2bsr(n, n, rp, ci, val, bs)
mb = (n + bs - 1) // bs
B = torch.zeros(mb * bs, K, device="cuda"); B[:n] = torch.rand(n, K, device="cuda") * 2 - 1
C = torch.zeros(mb * bs, K, device="cuda")
ts = []
for i in range(a.iters + 2):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    pkg.bsrmm(0, 1, mb, K, mb, bci.numel(), 1.0, bv, brp, bci, bs, B, K, 0.0, C, K, c_layout=0, math=a.math)
    e.record(); torch.cuda.synchronize()
    ts.append(s.elapsed_time(e))
ts = sorted(ts[2:])
nnzb = bci.numel()
print(json.dumps(dict(bs=bs, math=a.math, K=K, nnzb=nnzb, fill=ci.numel() / (nnzb * bs * bs), ms=ts[len(ts) // 2],
                      dense_tflops=2 * nnzb * bs * bs * K / ts[len(ts) // 2] / 1e9)))

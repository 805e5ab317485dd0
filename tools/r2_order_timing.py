#!/usr/bin/env python
"""Timings of the device orderings (spmmdb_order_bfs, spmmdb_order_by_degree) on the symmetrised config graphs, next
to the sequential restatement of the reference's BFSTraversal on the host (oracle, one core).  One JSON line each."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def symmetrise(rp, ci, n):
    import torch
    rows = torch.repeat_interleave(torch.arange(n, device=rp.device), (rp[1:] - rp[:-1]).long())
    a = torch.cat([rows, ci.long()])
    b = torch.cat([ci.long(), rows])
    key = torch.unique(a * n + b)
    a, b = key // n, key % n
    deg = torch.bincount(a, minlength=n)
    nrp = torch.zeros(n + 1, dtype=torch.int32, device=rp.device)
    nrp[1:] = torch.cumsum(deg, 0).to(torch.int32)
    return nrp, b.to(torch.int32)


def main():
    import torch

    import spmm_denseblock_b200 as pkg
    from oracle import oracle as O
    from spmm_denseblock_b200 import graphs
    pkg.lib()
    O.build()
    out = open(os.path.join(ROOT, "gpurun_out", "r2t_order_timing.jsonl"), "w")

    def timeit(fn, iters=5):
        fn()
        ts = []
        for _ in range(iters):
            torch.cuda.synchronize()
            t = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            ts.append((time.perf_counter() - t) * 1e3)
        ts.sort()
        return ts[len(ts) // 2]

    for cfg in ("c2_arxiv", "c4_molhiv", "c3_proteins", "c5_products"):
        for variant in ("as generated (directed)", "symmetrised"):
            rp, ci = graphs.make_config_graph(cfg, device="cuda")
            n = rp.numel() - 1
            if variant == "symmetrised":
                rp, ci = symmetrise(rp, ci, n)
            t_bfs = timeit(lambda: pkg.order_bfs(rp, ci))
            t_rcm = timeit(lambda: pkg.order_rcm(rp, ci))
            t_deg = timeit(lambda: pkg.order_by_degree(rp))
            p = pkg.order_bfs(rp, ci)
            t_apply = timeit(lambda: pkg.permute_csr(rp, ci, None, p))
            hrp, hci = rp.cpu().numpy(), ci.cpu().numpy()
            t = time.perf_counter()
            want = O.bfs_order(hrp, hci)
            t_cpu = (time.perf_counter() - t) * 1e3
            line = json.dumps(dict(config=cfg, pattern=variant, n=n, nnz=int(ci.numel()), order_bfs_ms=t_bfs,
                                   order_rcm_ms=t_rcm, order_by_degree_ms=t_deg, permute_csr_ms=t_apply,
                                   cpu_bfs_restatement_ms=t_cpu,
                                   bfs_equals_sequential=bool((p.cpu().numpy() == want).all()),
                                   rcm_equals_restatement=bool((pkg.order_rcm(rp, ci).cpu().numpy()
                                                                == O.rcm_order(hrp, hci)).all())))
            print(line, flush=True)
            out.write(line + "\n")
            del rp, ci, p
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

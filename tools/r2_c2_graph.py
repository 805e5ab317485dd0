#!/usr/bin/env python
"""C2 (ogbn-arxiv shape, 169,343 nodes, 1,166,243 edges, K = 128): the call is five launches on two streams for 28 us
of roofline time, so launch overhead matters ("time with CUDA graphs or back-to-back launches and
report both").  Times the drop-in call plain and replayed from a CUDA graph, on the small-matrix path and on the
single-kernel lean path, L2 flushed and warm.  One JSON line per case -> profiles/r2_c2_graph.jsonl."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "r2_c2_graph.jsonl")
    import spmm_denseblock_b200 as pkg
    from spmm_denseblock_b200 import graphs
    pkg.lib()
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] \
        if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out = open(out_path, "w")
    rp, ci = graphs.make_config_graph("c2_arxiv", device="cuda")
    n, nnz = rp.numel() - 1, ci.numel()
    val = torch.ones(nnz, device="cuda")

    def time_it(fn, do_flush, iters=30):
        for _ in range(5):
            fn()
        ts = []
        for _ in range(iters):
            if do_flush:
                flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            e.synchronize()
            ts.append(s.elapsed_time(e))
        ts.sort()
        return ts[len(ts) // 2], ts[0]

    for K in (64, 128, 256):
        g = torch.Generator(device="cuda").manual_seed(1234)
        B = torch.rand(n, K, generator=g, device="cuda") * 2 - 1
        C = torch.empty(n, K, device="cuda")
        Cref = torch.empty(n, K, device="cuda")
        pkg.csrmm(rp, ci, val, B, Cref, n, n, K)
        alg = 4 * (n + 1) + 8 * nnz + 8 * n * K
        for path, knobs in (("small-matrix path (scan + rows-pipe | long + reduce on a side stream)", {}),
                            ("lean kernel, inline segments of 64", {"CSR_SMALL": 0, "CSR_INLINE": 1}),
                            ("lean kernel, inline segments of 256", {"CSR_SMALL": 0, "CSR_INLINE": 1, "seg": 256})):
            for k_, v_ in knobs.items():
                if k_ == "seg":
                    pkg.set_csrmm_segment(v_)
                else:
                    pkg.set_tuning(k_, v_)
            call = lambda: pkg.csrmm(rp, ci, val, B, C, n, n, K, nnz=nnz)
            call()
            cap = torch.cuda.Stream()
            with torch.cuda.stream(cap):  # library scratch is per (device, stream): warm up on the capture stream
                call()
            torch.cuda.synchronize()
            ok = bool(torch.allclose(C, Cref, rtol=1e-4, atol=1e-4))
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=cap):
                call()
            for mode, fn in (("plain call", call), ("CUDA graph replay", gr.replay)):
                for fl in (True, False):
                    med, mn = time_it(fn, fl)
                    d = dict(config="c2_arxiv", n=n, nnz=nnz, K=K, path=path, launch=mode, l2="flushed" if fl else "warm",
                             ms_med=med, ms_min=mn, gflops=2.0 * nnz * K / med / 1e6, hbm_frac=alg / med / 1e6 / peak,
                             matches=ok)
                    print(json.dumps(d), flush=True)
                    out.write(json.dumps(d) + "\n")
            pkg.set_csrmm_segment(0)
            for k_ in knobs:
                if k_ != "seg":
                    pkg.set_tuning(k_, None)
    out.close()


if __name__ == "__main__":
    main()

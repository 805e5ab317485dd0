#!/usr/bin/env python
"""First contact with the tile::gather4 CSR kernel (CSR_TMA knob): bits against the lean kernel on a small hub matrix
and on C5 at scale 0.1, then timings on the full C5 graphs."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import spmm_denseblock_b200 as pkg  # noqa: E402
from spmm_denseblock_b200 import graphs  # noqa: E402


def timeit(fn, iters=8):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); e.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    stage = sys.argv[1] if len(sys.argv) > 1 else "small"
    pkg.lib()
    if stage == "small":
        from helpers import random_csr
        rng = np.random.default_rng(1)
        m, k = 6000, 9000
        rp, ci, v = random_csr(rng, m, k, 8, hubs={17: 5000, 1000: 8999, 2500: 513}, empty_rows={5, 6, 7})
        d = [torch.from_numpy(x).cuda() for x in (rp, ci, v)]
        pkg.set_tuning("CSR_SMALL", 0); pkg.set_tuning("CSR_INLINE", 1); pkg.set_csrmm_segment(256)
        for K in (256, 128, 64):
            B = torch.rand(k, K, device="cuda") * 2 - 1
            ref = torch.empty(m, K, device="cuda")
            pkg.csrmm(*d, B, ref, m, k, K)
            for mode in (1, 2):
                pkg.set_tuning("CSR_TMA", mode)
                out = torch.full((m, K), float("nan"), device="cuda")
                pkg.csrmm(*d, B, out, m, k, K)
                torch.cuda.synchronize()
                print(json.dumps(dict(stage="small fp32", K=K, mode=mode, equal=bool(torch.equal(out, ref)),
                                      maxdiff=float((out - ref).abs().nan_to_num(1e9).max()))), flush=True)
                Bb = pkg.dense_pack_bf16(B)
                pkg.set_tuning("CSR_TMA", 0)
                refb = torch.empty(m, K, device="cuda")
                pkg.csrmm_ex(*d, Bb, refb, m, k, K)
                pkg.set_tuning("CSR_TMA", mode)
                outb = torch.full((m, K), float("nan"), device="cuda")
                pkg.csrmm_ex(*d, Bb, outb, m, k, K)
                torch.cuda.synchronize()
                print(json.dumps(dict(stage="small bf16", K=K, mode=mode, equal=bool(torch.equal(outb, refb)))), flush=True)
                pkg.set_tuning("CSR_TMA", 0)
        return
    K = 256
    for gname, over in (("spec", {}), ("local", dict(p_local=0.9, window=4096))):
        rp, ci = graphs.make_config_graph("c5_products", device="cuda", **over)
        n, nnz = rp.numel() - 1, ci.numel()
        val = torch.ones(nnz, device="cuda")
        B = torch.rand(n, K, device="cuda") * 2 - 1
        Bb = pkg.dense_pack_bf16(B)
        ref, out = torch.empty(n, K, device="cuda"), torch.empty(n, K, device="cuda")
        for dt, call in (("fp32", lambda o: pkg.csrmm(rp, ci, val, B, o, n, n, K)),
                         ("bf16", lambda o: pkg.csrmm_ex(rp, ci, val, Bb, o, n, n, K))):
            pkg.set_tuning("CSR_TMA", 0)
            t0 = timeit(lambda: call(ref))
            print(json.dumps(dict(graph=gname, dtype=dt, kernel="lean", ms=t0)), flush=True)
            for mode in (1, 2):
                for rpi in (8, 16):
                    pkg.set_tuning("CSR_TMA", mode); pkg.set_tuning("CSR_RPI", rpi)
                    out.fill_(float("nan"))
                    t1 = timeit(lambda: call(out))
                    print(json.dumps(dict(graph=gname, dtype=dt, kernel=f"tma mode {mode} rpi {rpi}", ms=t1,
                                          equal=bool(torch.equal(out, ref)))), flush=True)
            pkg.set_tuning("CSR_TMA", 0); pkg.set_tuning("CSR_RPI", None)
        del rp, ci, val, B, Bb, ref, out
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

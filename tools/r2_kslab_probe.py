#!/usr/bin/env python
"""Does cutting K into column slabs (one csrmm call per slab, B and C addressed with ld = K) lower the DRAM traffic of
the products-shaped SpMM?  A slab of Ks columns makes a B row Ks * 4 bytes, so the 126 MB L2 holds K / Ks times more
rows of B.  Prints one JSON line per (graph, slab width)."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch

    import spmm_denseblock_b200 as pkg
    from spmm_denseblock_b200 import graphs
    L = pkg.lib()
    vp = C.c_void_p
    K = 256
    out = open(os.path.join(ROOT, "gpurun_out", "r2r_kslab_probe.jsonl"), "w")
    for gname, over in (("spec", {}), ("local", dict(p_local=0.9, window=4096))):
        rp, ci = graphs.make_config_graph("c5_products", device="cuda", **over)
        n, nnz = rp.numel() - 1, ci.numel()
        g = torch.Generator(device="cuda").manual_seed(1234)
        B = torch.rand(n, K, generator=g, device="cuda") * 2 - 1
        val = torch.ones(nnz, device="cuda")
        Cref = torch.empty(n, K, device="cuda")
        Cm = torch.empty(n, K, device="cuda")
        pkg.csrmm(rp, ci, val, B, Cref, n, n, K)
        for Ks in (256, 128, 64, 32):
            def run():
                for c0 in range(0, K, Ks):
                    rc = L.spmmdb_csrmm_f32(n, n, Ks, nnz, vp(rp.data_ptr()), vp(ci.data_ptr()), vp(val.data_ptr()),
                                            vp(B.data_ptr() + 4 * c0), K, 0, vp(Cm.data_ptr() + 4 * c0), K, 0, 1.0, 0.0,
                                            vp(torch.cuda.current_stream().cuda_stream))
                    assert rc == 0
            pkg.set_csrmm_segment(pkg.csrmm_segment(nnz, m=n))
            try:
                for _ in range(2):
                    run()
                ts = []
                for _ in range(6):
                    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    s.record()
                    run()
                    e.record()
                    e.synchronize()
                    ts.append(s.elapsed_time(e))
            finally:
                pkg.set_csrmm_segment(0)
            ts.sort()
            line = json.dumps(dict(graph=gname, n=n, nnz=nnz, K=K, slab_columns=Ks, calls=K // Ks, ms_med=ts[len(ts) // 2],
                                   ms_min=ts[0], bits_equal_one_call=bool(torch.equal(Cm, Cref))))
            print(line, flush=True)
            out.write(line + "\n")
        del B, val, Cref, Cm, rp, ci
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

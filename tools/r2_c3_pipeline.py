#!/usr/bin/env python
"""BASELINE.json config 3 the way north_star states it: the proteins-shaped planted-community graph with SHUFFLED
node ids -> CPU reordering by the reference's own code (reverseCuthillMcKee, vendored Gorder; permutations from
tests/golden/c3_orders.npz, made by tools/make_c3_orders.py) -> spmmdb_permute_csr_f32 + spmmdb_permute_rows_f32 ->
{CSR, csr2bsr + BSR tf32 / bf16 at bs 16 / 32, hybrid split + CSR + BSR} -> C un-permuted and compared with the
fp64 oracle on the ORIGINAL (shuffled) matrix.  One JSON line per (ordering, op): profiles/r2_c3_pipeline.jsonl.

    python tools/r2_c3_pipeline.py --out gpurun_out/r2_c3_pipeline.jsonl
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r2_c3_pipeline.jsonl"))
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--orders", default="shuffled,rcm,gorder,ideal")
    ap.add_argument("--densities", default="0.03,0.06,0.1,0.2")
    a = ap.parse_args()
    import spmm_denseblock_b200 as pkg
    from spmm_denseblock_b200 import graphs
    from oracle import oracle as O
    O.build()
    pkg.lib()
    dev = torch.device("cuda", 0)
    K = 128
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] \
        if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = open(a.out, "w")

    def emit(**kw):
        line = json.dumps(kw)
        print(line, flush=True)
        out.write(line + "\n")
        out.flush()

    def timeit(fn, warm=2):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(a.iters):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            e.synchronize()
            ts.append(s.elapsed_time(e))
        ts.sort()
        return ts[len(ts) // 2]

    t0 = time.time()
    rp_c, ci_c = graphs.make_config_graph("c3_proteins", shuffle=True)  # CPU generator: the fixtures' graph
    n, nnz = rp_c.numel() - 1, ci_c.numel()
    fx = np.load(os.path.join(ROOT, "tests", "golden", "c3_orders.npz"))
    fp = np.array([n, nnz, int(rp_c.numpy()[::997].astype(np.int64).sum() % (1 << 31)),
                   int((ci_c.numpy()[::9973].astype(np.int64) * 31 % 1000003).sum() % (1 << 31))], np.int64)
    assert np.array_equal(fp, fx["fingerprint"]), "tests/golden/c3_orders.npz belongs to another graph"
    rp0, ci0 = rp_c.to(dev), ci_c.to(dev)
    val0 = torch.ones(nnz, device=dev)
    g = torch.Generator(device=dev).manual_seed(1234)
    B0 = torch.rand(n, K, generator=g, device=dev) * 2 - 1
    # the oracle on the ORIGINAL matrix (fp64 accumulate, gespmm_csrmm.h:95-166 semantics) and its error scale
    ref = torch.from_numpy(O.csrmm(rp_c.numpy(), ci_c.numpy(), np.ones(nnz, np.float32), B0.cpu().numpy(), K)).to(dev)
    scale = torch.from_numpy(O.csrmm(rp_c.numpy(), ci_c.numpy(), np.ones(nnz, np.float32),
                                     B0.abs().cpu().numpy(), K)).to(dev).clamp_min(1e-30)
    emit(graph="c3_proteins shuffled", n=n, nnz=nnz, K=K, setup_s=time.time() - t0, hbm_peak_gbs=peak)
    # the planted order = the shuffle undone: an upper bound for any reordering
    rp_i, ci_i = graphs.make_config_graph("c3_proteins", shuffle=False)
    densities = [float(x) for x in a.densities.split(",")]

    for order in a.orders.split(","):
        if order == "ideal":
            # not a permutation of the shuffled graph's B (the planted graph is generated separately): its own oracle
            rp, ci = rp_i.to(dev), ci_i.to(dev)
            val, B = torch.ones(ci.numel(), device=dev), B0
            refo = torch.from_numpy(O.csrmm(rp_i.numpy(), ci_i.numpy(), np.ones(ci.numel(), np.float32),
                                            B0.cpu().numpy(), K)).to(dev)
            sco = torch.from_numpy(O.csrmm(rp_i.numpy(), ci_i.numpy(), np.ones(ci.numel(), np.float32),
                                           B0.abs().cpu().numpy(), K)).to(dev).clamp_min(1e-30)
            unperm = None
            t_perm = 0.0
        elif order == "shuffled":
            rp, ci, val, B, refo, sco, unperm, t_perm = rp0, ci0, val0, B0, ref, scale, None, 0.0
        else:
            old2new = torch.from_numpy(fx["old2new_" + order].astype(np.int32)).to(dev)
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            rp, ci, val = pkg.permute_csr(rp0, ci0, val0, old2new)
            B = pkg.permute_rows(B0, old2new)
            torch.cuda.synchronize()
            t_perm = (time.perf_counter() - t1) * 1e3
            new2old = torch.empty_like(old2new)
            new2old[old2new.long()] = torch.arange(n, device=dev, dtype=torch.int32)
            unperm, refo, sco = new2old, ref, scale
        m = rp.numel() - 1
        nz = ci.numel()

        def err_of(Cout):
            got = Cout[:m].contiguous()
            if unperm is not None:
                got = pkg.permute_rows(got, unperm)
            return float(((got.double() - refo).abs() / sco).max())

        base = dict(order=order, nnz=nz, permute_ms_once=t_perm)
        C = torch.empty(m, K, device=dev)
        ms = timeit(lambda: pkg.csrmm(rp, ci, val, B, C, m, m, K))
        ab = 4 * (m + 1) + 8 * nz + 8 * m * K
        emit(**base, op="csr fp32", ms=ms, useful_tflops=2.0 * nz * K / ms / 1e9, hbm_frac=ab / ms / 1e6 / peak,
             rel_err_vs_oracle=err_of(C))
        for bs in (16, 32):
            mb = (m + bs - 1) // bs
            mp = mb * bs
            Bp = torch.zeros(mp, K, device=dev)
            Bp[:m] = B
            hist = pkg.block_fill_histogram(rp, ci, bs)
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            brp, bci, bv = pkg.csr2bsr(m, m, rp, ci, val, bs)
            torch.cuda.synchronize()
            t_conv = (time.perf_counter() - t1) * 1e3
            nnzb = bci.numel()
            Cp = torch.empty(mp, K, device=dev)
            bbytes = 4 * (mb + 1) + 4 * nnzb + 4 * nnzb * bs * bs + 8 * mp * K
            for math, mname in ((pkg.MATH_TF32, "tf32"), (pkg.MATH_BF16, "bf16")):
                f = lambda: pkg.bsrmm(0, 1, mb, K, mb, nnzb, 1.0, bv, brp, bci, bs, Bp, K, 0.0, Cp, K,
                                      c_layout=pkg.ROW_MAJOR, math=math)
                ms = timeit(f)
                emit(**base, op=f"bsr{bs} {mname}", ms=ms, nnzb=nnzb, fill=nz / (nnzb * bs * bs), csr2bsr_ms_once=t_conv,
                     fill_deciles=hist, useful_tflops=2.0 * nz * K / ms / 1e9,
                     dense_tflops=2.0 * nnzb * bs * bs * K / ms / 1e9, hbm_frac=bbytes / ms / 1e6 / peak,
                     rel_err_vs_oracle=err_of(Cp))
            del brp, bci, bv
            for dens in densities:
                torch.cuda.synchronize()
                t1 = time.perf_counter()
                crp, cci, cv, brp, bci, bv = pkg.divide(m, m, rp, ci, val, bs, dens)
                torch.cuda.synchronize()
                t_div = (time.perf_counter() - t1) * 1e3
                nnzb = bci.numel()
                if nnzb == 0:
                    continue
                inb = nz - cci.numel()

                def hybrid():
                    pkg.csrmm(crp, cci, cv, Bp, Cp, m, m, K)
                    pkg.bsrmm(0, 1, mb, K, mb, nnzb, 1.0, bv, brp, bci, bs, Bp, K, 1.0, Cp, K, c_layout=pkg.ROW_MAJOR,
                              math=pkg.MATH_TF32)
                ms = timeit(hybrid)
                hb = 4 * (m + 1) + 8 * cci.numel() + 4 * (mb + 1) + 4 * nnzb + 4 * nnzb * bs * bs + 8 * mp * K
                emit(**base, op=f"hybrid{bs} tf32 density {dens}", ms=ms, nnzb=nnzb, nnz_in_blocks=inb,
                     share_in_blocks=inb / nz, block_fill=inb / (nnzb * bs * bs), remainder_nnz=cci.numel(),
                     divide_ms_once=t_div, useful_tflops=2.0 * nz * K / ms / 1e9, hbm_frac=hb / ms / 1e6 / peak,
                     rel_err_vs_oracle=err_of(Cp))
                del crp, cci, cv, brp, bci, bv
            del Bp, Cp
        torch.cuda.empty_cache()
    out.close()


if __name__ == "__main__":
    main()

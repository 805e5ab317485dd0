"""Developer timing sweep (not the bench contract): CUDA-event timings of every kernel
on the BASELINE.json config shapes, one JSON line per measurement."""
import argparse
import json
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import spmm_denseblock_b200 as pkg  # noqa: E402
from spmm_denseblock_b200 import graphs  # noqa: E402

HBM = 6546.9


def timeit(fn, warmup=3, iters=10, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            mode = os.environ.get("SPMMDB_FLUSH", "write")
            if mode == "write":        # L2 left full of DIRTY lines (their write-back lands inside the timed call)
                flush.add_(1.0)
            elif mode == "read":       # L2 left full of clean lines of another buffer
                flush.add_(1.0)
                flush.sum()
            # "warm": nothing -- the previous call's B / C stay in L2
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def csr_case(name, rp, ci, K, flush):
    n = rp.numel() - 1
    nnz = ci.numel()
    val = torch.ones(nnz, device="cuda")
    B = torch.rand(n, K, device="cuda") * 2 - 1
    C = torch.empty(n, K, device="cuda")
    med, best = timeit(lambda: pkg.csrmm(rp, ci, val, B, C, n, n, K, nnz=nnz), flush=flush)
    byts = 4 * (n + 1) + 8 * nnz + 4 * n * K * 2
    print(json.dumps(dict(case=name, kernel="csrmm", n=n, nnz=nnz, K=K, ms_med=med, ms_min=best,
                          gflops=2 * nnz * K / med / 1e6, gbs=byts / med / 1e6, frac=byts / med / 1e6 / HBM,
                          maxdeg=int((rp[1:] - rp[:-1]).max()))), flush=True)
    return val, B, C


def bsr_case(name, rp, ci, val, B, K, bs, maths, flush):
    n = rp.numel() - 1
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    med_c, _ = timeit(lambda: pkg.csr2bsr(n, n, rp, ci, val, bs), warmup=1, iters=3)
    brp, bci, bv = pkg.csr2bsr(n, n, rp, ci, val, bs)
    nnzb = bci.numel()
    nnz = ci.numel()
    mb = (n + bs - 1) // bs
    cb = 4 * (n + 1) + 8 * nnz + 4 * (mb + 1) + 4 * nnzb + 4 * nnzb * bs * bs
    print(json.dumps(dict(case=name, kernel="csr2bsr", bs=bs, nnzb=nnzb, fill=nnz / (nnzb * bs * bs), ms_med=med_c,
                          gbs=cb / med_c / 1e6, frac=cb / med_c / 1e6 / HBM)), flush=True)
    if nnzb * bs * bs < 2**31:
        med_b, _ = timeit(lambda: pkg.bsr2csr(mb, mb, brp, bci, bv, bs), warmup=1, iters=3)
        bb = 4 * (mb + 1) + 4 * nnzb + 4 * nnzb * bs * bs + 4 * (n + 1) + 8 * nnzb * bs * bs
        print(json.dumps(dict(case=name, kernel="bsr2csr", bs=bs, ms_med=med_b, gbs=bb / med_b / 1e6,
                              frac=bb / med_b / 1e6 / HBM)), flush=True)
    Bp = torch.zeros(mb * bs, K, device="cuda")
    Bp[:n] = B
    Cb = torch.zeros(mb * bs, K, device="cuda")
    bv16 = pkg.bsr_pack_bf16(bv, bs) if 3 in maths else None
    for math in maths:
        vals = bv16 if math == 3 else bv
        med, best = timeit(lambda: pkg.bsrmm(0, 1, mb, K, mb, nnzb, 1.0, vals, brp, bci, bs, Bp, K, 0.0, Cb, K,
                                             c_layout=0, math=math), warmup=2, iters=5, flush=flush)
        byts = 4 * (mb + 1) + 4 * nnzb + (2 if math == 3 else 4) * nnzb * bs * bs + 4 * n * K * 2
        print(json.dumps(dict(case=name, kernel="bsrmm", math=math, bs=bs, ms_med=med, ms_min=best,
                              useful_gflops=2 * nnz * K / med / 1e6, dense_tflops=2 * nnzb * bs * bs * K / med / 1e9,
                              gbs=byts / med / 1e6, frac_hbm=byts / med / 1e6 / HBM)), flush=True)


def hybrid_case(name, rp, ci, K, bs, densities, math, flush):
    """divide.cu's idea on the device: dense blocks -> tcgen05 BSR path, the rest -> CSR path, one C."""
    n, nnz = rp.numel() - 1, ci.numel()
    mb = (n + bs - 1) // bs
    val = torch.ones(nnz, device="cuda")
    Bp = torch.rand(mb * bs, K, device="cuda") * 2 - 1
    C = torch.zeros(mb * bs, K, device="cuda")
    med_csr, _ = timeit(lambda: pkg.csrmm(rp, ci, val, Bp, C, n, n, K, nnz=nnz), flush=flush)
    ref = C.clone()
    for dens in densities:
        med_d, _ = timeit(lambda: pkg.divide(n, n, rp, ci, None, bs, dens), warmup=1, iters=3)
        crp, cci, cv, brp, bci, bv = pkg.divide(n, n, rp, ci, None, bs, dens)
        nnzb, cnnz = bci.numel(), cci.numel()

        def run():
            pkg.csrmm(crp, cci, cv, Bp, C, n, n, K, nnz=cnnz)  # beta = 0: C overwritten
            if nnzb:
                pkg.bsrmm(0, 1, mb, K, mb, nnzb, 1.0, bv, brp, bci, bs, Bp, K, 1.0, C, K, c_layout=0, math=math)
        med, best = timeit(run, warmup=2, iters=7, flush=flush)
        err = float((C[:n] - ref[:n]).abs().max() / ref[:n].abs().max())
        print(json.dumps(dict(case=name, kernel="hybrid", bs=bs, density=dens, math=math, nnzb=nnzb, csr_nnz=cnnz,
                              dense_part=1 - cnnz / nnz, block_fill=(nnz - cnnz) / max(nnzb * bs * bs, 1),
                              ms_hybrid=med, ms_csr_only=med_csr, speedup=med_csr / med, ms_divide=med_d,
                              useful_gflops=2 * nnz * K / med / 1e6, rel_diff_vs_csr=err)), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="c2,c2l,c4,c5,c5l,c3")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--maths", default="0")
    a = ap.parse_args()
    flush = torch.empty(256 << 20, dtype=torch.float32, device="cuda")  # 1 GiB > L2
    maths = [int(x) for x in a.maths.split(",")]
    for case in a.cases.split(","):
        if case == "c2":
            rp, ci = graphs.make_config_graph("c2_arxiv", device="cuda", scale=a.scale)
            csr_case("c2_arxiv_random", rp, ci, 128, flush)
        elif case == "c2l":
            rp, ci = graphs.make_config_graph("c2_arxiv", device="cuda", scale=a.scale, p_local=0.8, window=512)
            csr_case("c2_arxiv_local", rp, ci, 128, flush)
        elif case == "c4csr":
            rp, ci = graphs.make_config_graph("c4_molhiv", device="cuda", scale=a.scale)
            csr_case("c4_molhiv", rp, ci, 300, flush)
        elif case == "c4":
            rp, ci = graphs.make_config_graph("c4_molhiv", device="cuda", scale=a.scale)
            val, B, C = csr_case("c4_molhiv", rp, ci, 300, flush)
            for bs in (16, 32):
                bsr_case("c4_molhiv", rp, ci, val, B, 300, bs, maths, flush)
        elif case == "c5":
            rp, ci = graphs.make_config_graph("c5_products", device="cuda", scale=a.scale)
            csr_case("c5_products_random", rp, ci, 256, flush)
        elif case == "c5l":
            rp, ci = graphs.make_config_graph("c5_products", device="cuda", scale=a.scale, p_local=0.9, window=4096)
            csr_case("c5_products_local", rp, ci, 256, flush)
        elif case == "c3csr":
            rp, ci = graphs.make_config_graph("c3_proteins", device="cuda", scale=a.scale)
            csr_case("c3_proteins_ideal", rp, ci, 128, flush)
        elif case == "c3d":
            # block-dense variant: every edge stays inside its community (what the dense part of a
            # hybrid split looks like)
            rp, ci = graphs.make_config_graph("c3_proteins", device="cuda", scale=a.scale, p_in=1.0)
            val, B, C = csr_case("c3_proteins_pin1", rp, ci, 128, flush)
            for bs in (16, 32):
                bsr_case("c3_proteins_pin1", rp, ci, val, B, 128, bs, maths, flush)
        elif case == "perm":
            # permutation application on the device (permutate, reorder_strategy.cc:26-39) + heat map
            for name, K in (("c5_products", 256), ("c3_proteins", 128), ("c2_arxiv", 128)):
                rp, ci = graphs.make_config_graph(name, device="cuda", scale=a.scale)
                n, nnz = rp.numel() - 1, ci.numel()
                perm = torch.randperm(n, device="cuda").to(torch.int32)
                val = torch.ones(nnz, device="cuda")
                X = torch.rand(n, K, device="cuda")
                m1, _ = timeit(lambda: pkg.permute_csr(rp, ci, val, perm), warmup=1, iters=3)
                m2, _ = timeit(lambda: pkg.permute_rows(X, perm), warmup=1, iters=3)
                m3, _ = timeit(lambda: pkg.heatmap(rp, ci, 256), warmup=1, iters=3)
                b1 = 8 * (n + 1) + 4 * n + 16 * nnz + 4 * nnz  # rowptr in/out, old2new, (col, val) in/out, old2new gathers
                print(json.dumps(dict(case=name, kernel="permute_csr", n=n, nnz=nnz, ms_med=m1, gbs=b1 / m1 / 1e6,
                                      frac=b1 / m1 / 1e6 / HBM, maxdeg=int((rp[1:] - rp[:-1]).max()),
                                      permute_rows_ms=m2, permute_rows_gbs=8 * n * K / m2 / 1e6, heatmap256_ms=m3)),
                      flush=True)
                del X, val
                torch.cuda.empty_cache()
        elif case == "c3h":
            # hybrid split on the community graph as generated (90 % of the edges inside communities) and on a
            # noisier one (60 %): how much of the matrix the dense path takes and what it buys
            for p_in in (0.9, 0.6):
                rp, ci = graphs.make_config_graph("c3_proteins", device="cuda", scale=a.scale, p_in=p_in)
                for bs in (16, 32):
                    hybrid_case(f"c3_proteins_pin{p_in}", rp, ci, 128, bs, (0.02, 0.05, 0.1, 0.2), 1, flush)
        elif case == "c3":
            rp, ci = graphs.make_config_graph("c3_proteins", device="cuda", scale=a.scale)
            val, B, C = csr_case("c3_proteins_ideal", rp, ci, 128, flush)
            for bs in (16, 32):
                bsr_case("c3_proteins_ideal", rp, ci, val, B, 128, bs, maths, flush)
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

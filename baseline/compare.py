#!/usr/bin/env python
"""ctypes bindings of the comparators (baseline/_ref/libref_gpu.so = the reference's own GPU kernels compiled
for sm_100a, baseline/_ref/libvendor_cmp.so = the cuSPARSE / cuBLAS calls the reference makes) and a timing
harness that runs them next to libspmmdb.so on the config-shaped inputs:

    python baseline/compare.py --out profiles/r2_vendor.jsonl [--configs c2,c3,c4,c5] [--scale 1.0]

Comparator / test infrastructure: nothing in spmm_denseblock_b200/ imports this.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

vp, i64 = C.c_void_p, C.c_int64
_REF = _VEN = None


def build():
    subprocess.check_call(["make", "-s", "-C", HERE])


def _load(name):
    p = os.path.join(HERE, "_ref", name)
    if not os.path.exists(p):
        build()
    return C.CDLL(p) if os.path.exists(p) else None


def ref_gpu():
    """The reference's own kernels; None when they could not be built (no /root/reference, no prebuilt .so)."""
    global _REF
    if _REF is None:
        L = _load("libref_gpu.so")
        if L is None:
            return None
        L.ref_gespmm_csrmm_f32.argtypes = [C.c_int, C.c_int, vp, vp, vp, vp, vp]
        L.ref_spmm_wrapper_f32.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp]
        L.ref_rocsparse_bsrmm_f32.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, vp, vp,
                                              vp, C.c_int, vp, C.c_int, C.c_float, vp, C.c_int]
        _REF = L
    return _REF


def vendor():
    global _VEN
    if _VEN is None:
        L = _load("libvendor_cmp.so")
        L.vendor_csr2bsr_nnz.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp, C.c_int, vp, C.POINTER(C.c_int)]
        L.vendor_csr2bsr.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int, vp, vp, vp]
        L.vendor_bsr2csr.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int, vp, vp, vp]
        L.vendor_bsrmm.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, vp, vp, vp, C.c_int,
                                   vp, C.c_int, C.c_float, vp, C.c_int]
        L.vendor_spmm_csr.argtypes = [C.c_int, C.c_int, C.c_int, i64, vp, vp, vp, vp, i64, C.c_int, vp, i64, C.c_int,
                                      C.c_float, C.c_float, C.c_int]
        L.vendor_block_gemm_loop.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp]
        _VEN = L
    return _VEN


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else C.c_void_p(0)


def _ok(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} failed: {rc}")


# ---- thin operators on torch CUDA tensors (default stream, like the reference) --------------------------------
def ref_gespmm(m, K, rp, ci, val, B, Cout):
    import torch
    with torch.cuda.stream(torch.cuda.default_stream()):
        _ok(ref_gpu().ref_gespmm_csrmm_f32(m, K, _p(rp), _p(ci), _p(val), _p(B), _p(Cout)), "ref gespmm_csrmm")
    return Cout


def ref_rocsparse_bsrmm(direction, trans_b, mb, n, kb, nnzb, alpha, bv, brp, bci, bs, B, ldb, beta, Cout, ldc):
    _ok(ref_gpu().ref_rocsparse_bsrmm_f32(direction, trans_b, mb, n, kb, nnzb, alpha, _p(bv), _p(brp), _p(bci), bs,
                                          _p(B), ldb, beta, _p(Cout), ldc), "ref rocsparse_bsrmm_template")
    return Cout


def cusparse_csr2bsr(m, n, rp, ci, val, bs, direction=0):
    import torch
    mb = (m + bs - 1) // bs
    brp = torch.empty(mb + 1, dtype=torch.int32, device=rp.device)
    nnzb = C.c_int(0)
    _ok(vendor().vendor_csr2bsr_nnz(direction, m, n, _p(rp), _p(ci), bs, _p(brp), C.byref(nnzb)), "cusparseXcsr2bsrNnz")
    bci = torch.empty(nnzb.value, dtype=torch.int32, device=rp.device)
    bv = torch.empty(nnzb.value * bs * bs, dtype=torch.float32, device=rp.device)
    _ok(vendor().vendor_csr2bsr(direction, m, n, _p(rp), _p(ci), _p(val), bs, _p(brp), _p(bci), _p(bv)),
        "cusparseScsr2bsr")
    return brp, bci, bv


def cusparse_bsr2csr(mb, nb, brp, bci, bv, bs, direction=0):
    import torch
    nnzb = bci.numel()
    rp = torch.empty(mb * bs + 1, dtype=torch.int32, device=brp.device)
    ci = torch.empty(nnzb * bs * bs, dtype=torch.int32, device=brp.device)
    v = torch.empty(nnzb * bs * bs, dtype=torch.float32, device=brp.device)
    _ok(vendor().vendor_bsr2csr(direction, mb, nb, _p(brp), _p(bci), _p(bv), bs, _p(rp), _p(ci), _p(v)),
        "cusparseSbsr2csr")
    return rp, ci, v


def cusparse_bsrmm(direction, trans_b, mb, n, kb, nnzb, alpha, bv, brp, bci, bs, B, ldb, beta, Cout, ldc):
    _ok(vendor().vendor_bsrmm(direction, trans_b, mb, n, kb, nnzb, alpha, _p(bv), _p(brp), _p(bci), bs, _p(B), ldb,
                              beta, _p(Cout), ldc), "cusparseSbsrmm")
    return Cout


def cusparse_spmm(m, k, K, rp, ci, val, B, Cout, alg=2, ldb=None, b_layout=0, ldc=None, c_layout=0, alpha=1.0, beta=0.0):
    ldb = (K if b_layout == 0 else k) if ldb is None else ldb
    ldc = (K if c_layout == 0 else m) if ldc is None else ldc
    _ok(vendor().vendor_spmm_csr(m, k, K, ci.numel(), _p(rp), _p(ci), _p(val), _p(B), ldb, b_layout, _p(Cout), ldc,
                                 c_layout, alpha, beta, alg), "cusparseSpMM")
    return Cout


def cublas_block_gemm_loop(mb, dim, bs, brp_host, bci_host, bv, Y, Z):
    """brp_host / bci_host: numpy int32 arrays (block_cublas.cu keeps the index arrays on the host)."""
    _ok(vendor().vendor_block_gemm_loop(mb, dim, bs, brp_host.ctypes.data_as(vp), bci_host.ctypes.data_as(vp), _p(bv),
                                        _p(Y), _p(Z)), "cublasSgemm loop")
    return Z


# ---- timing harness ---------------------------------------------------------------------------------------------
def _time(fn, warm=3, iters=10, flush=None):
    """Median / min ms of fn() with CUDA events on the default stream (the comparators launch there)."""
    import torch
    ts = []
    s = torch.cuda.default_stream()
    with torch.cuda.stream(s):
        for _ in range(warm):
            fn()
        for _ in range(iters):
            if flush is not None:
                flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(s)
            fn()
            b.record(s)
            b.synchronize()
            ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r2_vendor.jsonl"))
    ap.add_argument("--configs", default="c2,c3,c4,c5")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    import numpy as np
    import torch

    import spmm_denseblock_b200 as pkg
    from spmm_denseblock_b200 import graphs
    pkg.lib()
    torch.cuda.set_device(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] \
        if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    rows = []

    def emit(**kw):
        rows.append(kw)
        print(json.dumps(kw), flush=True)

    def maxrel(x, y):
        return float((x - y).abs().max() / max(float(y.abs().max()), 1e-30))

    cases = {"c2": ("c2_arxiv", {}, 128), "c3": ("c3_proteins", dict(p_in=0.9), 128),
             "c4": ("c4_molhiv", {}, 300), "c5": ("c5_products", {}, 256)}
    for key in a.configs.split(","):
        name, over, K = cases[key]
        rp, ci = graphs.make_config_graph(name, device="cuda", scale=a.scale, **over)
        n, nnz = rp.numel() - 1, ci.numel()
        g = torch.Generator(device="cuda").manual_seed(1234)
        B = torch.rand(n, K, generator=g, device="cuda") * 2 - 1
        val = torch.ones(nnz, device="cuda")
        Cm = torch.empty(n, K, device="cuda")
        Cr = torch.empty(n, K, device="cuda")
        alg_bytes = 4 * (n + 1) + 8 * nnz + 8 * n * K
        base = dict(config=key, n=n, nnz=nnz, K=K, op="csrmm", flops=2.0 * nnz * K, alg_bytes=alg_bytes)
        with torch.cuda.stream(torch.cuda.default_stream()):
            ours = lambda: pkg.csrmm(rp, ci, val, B, Cm, n, n, K)
            med, mn = _time(ours, iters=a.iters, flush=flush)
            emit(**base, impl="spmmdb_csrmm_f32", ms_med=med, ms_min=mn, gflops=2.0 * nnz * K / med / 1e6,
                 hbm_frac=alg_bytes / med / 1e6 / peak)
            if ref_gpu() is not None:
                med, mn = _time(lambda: ref_gespmm(n, K, rp, ci, val, B, Cr), iters=a.iters, flush=flush)
                emit(**base, impl="reference gespmm_csrmm (spmm_test2, sm_100a build)", ms_med=med, ms_min=mn,
                     gflops=2.0 * nnz * K / med / 1e6, hbm_frac=alg_bytes / med / 1e6 / peak,
                     max_rel_diff_vs_ours=maxrel(Cr, Cm), bit_equal_rows=float((Cr == Cm).all(1).float().mean()))
            for alg in (2, 3) if K % 4 == 0 else (2,):
                try:
                    med, mn = _time(lambda: cusparse_spmm(n, n, K, rp, ci, val, B, Cr, alg=alg), iters=a.iters,
                                    flush=flush)
                    emit(**base, impl=f"cusparseSpMM CSR_ALG{alg} (preprocessed)", ms_med=med, ms_min=mn,
                         gflops=2.0 * nnz * K / med / 1e6, hbm_frac=alg_bytes / med / 1e6 / peak,
                         max_rel_diff_vs_ours=maxrel(Cr, Cm))
                except RuntimeError as e:
                    emit(**base, impl=f"cusparseSpMM CSR_ALG{alg}", error=str(e))
            del Cr
            if key in ("c3", "c4"):
                for bs in (16, 32):
                    mb = (n + bs - 1) // bs
                    t_ours = _time(lambda: pkg.csr2bsr(n, n, rp, ci, val, bs), warm=1, iters=5, flush=flush)
                    brp, bci, bv = pkg.csr2bsr(n, n, rp, ci, val, bs)
                    nnzb = bci.numel()
                    cb = 4 * (n + 1) + 8 * nnz + 4 * (mb + 1) + 4 * nnzb + 4 * nnzb * bs * bs
                    emit(config=key, op="csr2bsr", bs=bs, nnzb=nnzb, fill=nnz / (nnzb * bs * bs), impl="spmmdb_csr2bsr",
                         ms_med=t_ours[0], ms_min=t_ours[1], alg_bytes=cb, hbm_frac=cb / t_ours[0] / 1e6 / peak)
                    # cuSPARSE indexes bsr_val with 32-bit ints: nnzb * bs^2 >= 2^31 is an illegal access inside the
                    # library (seen on C3 bs 32: 4.02 M blocks x 1024), after which the context is gone
                    fits32 = nnzb * bs * bs < 2 ** 31
                    if fits32:
                        t_v = _time(lambda: cusparse_csr2bsr(n, n, rp, ci, val, bs), warm=1, iters=5, flush=flush)
                        vrp, vci, vv = cusparse_csr2bsr(n, n, rp, ci, val, bs)
                        emit(config=key, op="csr2bsr", bs=bs, nnzb=nnzb, impl="cusparseXcsr2bsrNnz + cusparseScsr2bsr",
                             ms_med=t_v[0], ms_min=t_v[1], alg_bytes=cb, hbm_frac=cb / t_v[0] / 1e6 / peak,
                             bit_equal=bool(torch.equal(vrp, brp) and torch.equal(vci, bci) and torch.equal(vv, bv)))
                        del vrp, vci, vv
                    else:
                        emit(config=key, op="csr2bsr", bs=bs, nnzb=nnzb, impl="cusparseXcsr2bsrNnz + cusparseScsr2bsr",
                             error="skipped: nnzb * bs^2 >= 2^31 overflows cuSPARSE's 32-bit value index")
                    if fits32 and nnzb * bs * bs * 12 < 20e9:
                        bb = 4 * (mb + 1) + 4 * nnzb + 4 * nnzb * bs * bs + 4 * (n + 1) + 8 * nnzb * bs * bs
                        t_o = _time(lambda: pkg.bsr2csr(mb, mb, brp, bci, bv, bs), warm=1, iters=5, flush=flush)
                        o = pkg.bsr2csr(mb, mb, brp, bci, bv, bs)
                        emit(config=key, op="bsr2csr", bs=bs, impl="spmmdb_bsr2csr", ms_med=t_o[0], ms_min=t_o[1],
                             alg_bytes=bb, hbm_frac=bb / t_o[0] / 1e6 / peak)
                        t_c = _time(lambda: cusparse_bsr2csr(mb, mb, brp, bci, bv, bs), warm=1, iters=5, flush=flush)
                        c = cusparse_bsr2csr(mb, mb, brp, bci, bv, bs)
                        emit(config=key, op="bsr2csr", bs=bs, impl="cusparseSbsr2csr", ms_med=t_c[0], ms_min=t_c[1],
                             alg_bytes=bb, hbm_frac=bb / t_c[0] / 1e6 / peak,
                             bit_equal=bool(all(torch.equal(x, y) for x, y in zip(o, c))))
                        del o, c
                    # BSR SpMM: C column-major like the reference (run_bsrmm.cu:158-164), B row-major (trans_b = T)
                    mp = mb * bs
                    Bp = torch.zeros(mp, K, device="cuda")
                    Bp[:n] = B
                    Z = torch.zeros(K, mp, device="cuda")
                    Zv = torch.zeros(K, mp, device="cuda")
                    dense = 2.0 * nnzb * bs * bs * K
                    bbytes = 4 * (mb + 1) + 4 * nnzb + 4 * nnzb * bs * bs + 4 * mp * K + 4 * mp * K
                    bbase = dict(config=key, op="bsrmm", bs=bs, nnzb=nnzb, fill=nnz / (nnzb * bs * bs), K=K,
                                 dense_flops=dense, alg_bytes=bbytes)
                    if fits32:
                        med, mn = _time(lambda: cusparse_bsrmm(0, 1, mb, K, mb, nnzb, 1.0, bv, brp, bci, bs, Bp, K, 0.0,
                                                               Zv, mp), iters=a.iters, flush=flush)
                        emit(**bbase, impl="cusparseSbsrmm", ms_med=med, ms_min=mn, dense_tflops=dense / med / 1e9,
                             useful_tflops=2.0 * nnz * K / med / 1e9, hbm_frac=bbytes / med / 1e6 / peak)
                    else:
                        pkg.bsrmm(0, 1, mb, K, mb, nnzb, 1.0, bv, brp, bci, bs, Bp, K, 0.0, Zv, mp, math=pkg.MATH_FP32)
                    for math, mname in ((pkg.MATH_FP32, "fp32 CUDA cores"), (pkg.MATH_TF32, "tf32 tcgen05"),
                                        (pkg.MATH_BF16, "bf16 tcgen05")):
                        f = lambda: pkg.bsrmm(0, 1, mb, K, mb, nnzb, 1.0, bv, brp, bci, bs, Bp, K, 0.0, Z, mp, math=math)
                        med, mn = _time(f, iters=a.iters, flush=flush)
                        emit(**bbase, impl=f"spmmdb_bsrmm_f32 {mname}", ms_med=med, ms_min=mn,
                             dense_tflops=dense / med / 1e9, useful_tflops=2.0 * nnz * K / med / 1e9,
                             hbm_frac=bbytes / med / 1e6 / peak, max_rel_diff_vs_cusparse_or_fp32=maxrel(Z, Zv))
                    if ref_gpu() is not None and fits32:
                        Zr = torch.full((K, mp), float("nan"), device="cuda")
                        med, mn = _time(lambda: ref_rocsparse_bsrmm(0, 1, mb, K, mb, nnzb, 1.0, bv, brp, bci, bs, Bp, K,
                                                                    0.0, Zr, mp), iters=a.iters, flush=flush)
                        emit(**bbase, impl="reference rocsparse_bsrmm_template (sm_100a build)", ms_med=med, ms_min=mn,
                             dense_tflops=dense / med / 1e9, entries_written=float((~torch.isnan(Zr)).float().mean()),
                             note="output-column bug rocsparse_bsrmm_impl.h:326: C is not written for n <= 256")
                        del Zr
                    # block-GEMM formulation (block_cublas.cu:123-136): column-major blocks
                    if key == "c4":
                        bvc = bv.view(nnzb, bs, bs).transpose(1, 2).contiguous().view(-1)
                        Zb = torch.zeros(K, mp, device="cuda")
                        for math, mname in ((pkg.MATH_FP32, "fp32"), (pkg.MATH_TF32, "tf32")):
                            f = lambda: pkg.block_gemm(mb, K, nnzb, bs, brp, bci, bvc, Bp, Zb, math=math)
                            med, mn = _time(f, iters=a.iters, flush=flush)
                            emit(**bbase, impl=f"spmmdb_block_gemm_f32 {mname} (one grouped launch)", ms_med=med,
                                 ms_min=mn, dense_tflops=dense / med / 1e9)
                        sample = min(mb, 4096)  # the cuBLAS loop is launch bound: time a prefix of the block rows
                        hrp = brp[:sample + 1].cpu().numpy().astype(np.int32)
                        hci = bci.cpu().numpy().astype(np.int32)
                        Zc = torch.zeros(K, mp, device="cuda")
                        med, mn = _time(lambda: cublas_block_gemm_loop(sample, K, bs, hrp, hci, bvc, Bp, Zc), warm=1,
                                        iters=3)
                        nb_s = int(hrp[-1])
                        emit(**bbase, impl="cublasSgemm loop (block_cublas.cu:123-136)", ms_med=med * nnzb / max(nb_s, 1),
                             sample_blocks=nb_s, sample_ms=med, note="time scaled from a prefix of the block rows",
                             dense_tflops=dense / (med * nnzb / max(nb_s, 1)) / 1e9)
                    del brp, bci, bv, Bp, Z, Zv
        del rp, ci, B, val, Cm
        torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as f:
        for r in rows:
            f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()

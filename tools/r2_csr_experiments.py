#!/usr/bin/env python
"""Round-2 CSR experiments on the products-shaped graph (C5): fp32 / bf16 B, hot-column L2 policies, kernel
variants.  Writes one JSON line per measurement (profiles/r2_csr_experiments.jsonl is a copy of a run).

    python tools/r2_csr_experiments.py --out gpurun_out/r2_csr_experiments.jsonl [--graphs spec,local] [--quick]
"""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r2_csr_experiments.jsonl"))
    ap.add_argument("--graphs", default="spec,local")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--iters", type=int, default=8)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--only", default="")
    ap.add_argument("--policies", type=int, default=1, help="0: no L2-policy cases, 1: the short list, 2: the r2a sweep")
    a = ap.parse_args()
    import torch

    import spmm_denseblock_b200 as pkg
    from spmm_denseblock_b200 import graphs
    pkg.lib()
    torch.cuda.set_device(0)
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] \
        if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    rt = C.CDLL("libcudart.so.12")
    out = open(a.out, "w")

    def emit(**kw):
        line = json.dumps(kw)
        print(line, flush=True)
        out.write(line + "\n")
        out.flush()

    def timeit(fn, warm=3):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(a.iters):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            e.synchronize()
            ts.append(s.elapsed_time(e))
        ts.sort()
        return ts[len(ts) // 2], ts[0]

    props = torch.cuda.get_device_properties(0)
    l2 = props.L2_cache_size
    max_persist = C.c_int(0)
    rt.cudaDeviceGetAttribute(C.byref(max_persist), 108, 0)  # cudaDevAttrMaxPersistingL2CacheSize
    emit(device=props.name, l2_bytes=l2, max_persisting_l2=max_persist.value, hbm_peak_gbs=peak)

    K = 256
    for gname in a.graphs.split(","):
        over = dict(p_local=0.9, window=4096) if gname == "local" else {}
        rp, ci = graphs.make_config_graph("c5_products", device="cuda", scale=a.scale, **over)
        n, nnz = rp.numel() - 1, ci.numel()
        g = torch.Generator(device="cuda").manual_seed(1234)
        B = torch.rand(n, K, generator=g, device="cuda") * 2 - 1
        Bb = pkg.dense_pack_bf16(B)
        val = torch.ones(nnz, device="cuda")
        Cm = torch.empty(n, K, device="cuda")
        Cref = torch.empty(n, K, device="cuda")
        alg = 4 * (n + 1) + 8 * nnz + 8 * n * K
        alg_b = 4 * (n + 1) + 8 * nnz + 6 * n * K
        base = dict(graph=gname, n=n, nnz=nnz, K=K, maxdeg=int((rp[1:] - rp[:-1]).max()))

        def run(label, fn, bytes_alg, **kw):
            if a.only and a.only not in label:
                return None
            med, mn = timeit(fn)
            emit(**base, case=label, ms_med=med, ms_min=mn, gflops=2.0 * nnz * K / med / 1e6,
                 alg_gbs=bytes_alg / med / 1e6, hbm_frac=bytes_alg / med / 1e6 / peak,
                 gather_tbs=nnz * K * (2 if "bf16" in label else 4) / med / 1e9, **kw)
            return med

        run("fp32 drop-in", lambda: pkg.csrmm(rp, ci, val, B, Cref, n, n, K), alg)
        for bulk in (1,):
            pkg.set_tuning("CSR_LEAN", bulk)
            for rpi in (8,):
                pkg.set_tuning("CSR_RPI", rpi)
                Cm.fill_(float("nan"))
                run(f"fp32 bulk {bulk} rpi {rpi}", lambda: pkg.csrmm(rp, ci, val, B, Cm, n, n, K), alg,
                    bit_equal=None)
                emit(**base, case=f"fp32 bulk {bulk} rpi {rpi} bits equal to the register kernel",
                     value=bool(torch.equal(Cm, Cref)))
            pkg.set_tuning("CSR_RPI", None)
            pkg.set_tuning("CSR_LEAN", None)
        if a.policies:
            hot, n_hot = pkg.csr_hot_columns(n, ci, 64 * (1 << 20) // (K * 4))
            pkg.set_tuning("CSR_HOT_POLICY", 1)
            for pol in (1, 0, 2):
                pkg.set_tuning("CSR_COLD_POLICY", pol)
                run(f"fp32 hot(evict_normal hint) 64 MB cold_policy {pol}",
                    lambda: pkg.csrmm_ex(rp, ci, val, B, Cm, n, n, K, hot=hot), alg, n_hot=n_hot)
            pkg.set_tuning("CSR_COLD_POLICY", None)
            pkg.set_tuning("CSR_HOT_POLICY", None)
        for variant in ((2, 3) if not a.quick and a.policies else ()):
            pkg.set_tuning("CSR_VARIANT", variant)
            run(f"fp32 variant {variant}", lambda: pkg.csrmm(rp, ci, val, B, Cm, n, n, K), alg)
            pkg.set_tuning("CSR_VARIANT", None)
        for budget_mb in ((48, 64, 80, 96) if not a.quick else (64,)) if a.policies == 2 else ():
            rows = budget_mb * (1 << 20) // (K * 4)
            hot, n_hot = pkg.csr_hot_columns(n, ci, rows)
            for pol in (1, 0, 2) if not a.quick else (1,):
                pkg.set_tuning("CSR_COLD_POLICY", pol)
                run(f"fp32 hot {budget_mb} MB cold_policy {pol}",
                    lambda: pkg.csrmm_ex(rp, ci, val, B, Cm, n, n, K, hot=hot), alg, n_hot=n_hot)
                pkg.set_tuning("CSR_COLD_POLICY", None)
            assert torch.equal(Cm, Cref)
        # with an L2 set-aside for persisting (evict_last) lines
        if max_persist.value > 0 and not a.quick and a.policies == 2:
            for frac in (0.5, 0.75):
                sz = int(max_persist.value * frac)
                rt.cudaDeviceSetLimit(6, C.c_size_t(sz))  # cudaLimitPersistingL2CacheSize
                rows = int(sz * 0.9) // (K * 4)
                hot, n_hot = pkg.csr_hot_columns(n, ci, rows)
                pkg.set_tuning("CSR_COLD_POLICY", 1)
                run(f"fp32 hot set-aside {sz >> 20} MB", lambda: pkg.csrmm_ex(rp, ci, val, B, Cm, n, n, K, hot=hot), alg,
                    n_hot=n_hot)
                run(f"fp32 drop-in with set-aside {sz >> 20} MB", lambda: pkg.csrmm(rp, ci, val, B, Cm, n, n, K), alg)
                pkg.set_tuning("CSR_COLD_POLICY", None)
            rt.cudaDeviceSetLimit(6, C.c_size_t(0))
        # bf16-stored B
        pkg.set_tuning("CSR_LEAN", 0)
        for variant in (1,):
            pkg.set_tuning("CSR_VARIANT", variant)
            run(f"bf16 B variant {variant}", lambda: pkg.csrmm_ex(rp, ci, val, Bb, Cm, n, n, K), alg_b)
            pkg.set_tuning("CSR_VARIANT", None)
        pkg.set_tuning("CSR_LEAN", None)
        err = float((Cm - Cref).abs().max() / Cref.abs().max())
        emit(**base, case="bf16 B max |diff| / max |C| vs fp32", value=err)
        Cb = Cm.clone()
        for bulk in (1, 3):
            pkg.set_tuning("CSR_LEAN", bulk)
            for rpi in (8, 16):
                pkg.set_tuning("CSR_RPI", rpi)
                Cm.fill_(float("nan"))
                run(f"bf16 B bulk {bulk} rpi {rpi}", lambda: pkg.csrmm_ex(rp, ci, val, Bb, Cm, n, n, K), alg_b)
                emit(**base, case=f"bf16 B bulk {bulk} rpi {rpi} bits equal to the register kernel",
                     value=bool(torch.equal(Cm, Cb)))
            pkg.set_tuning("CSR_RPI", None)
            pkg.set_tuning("CSR_LEAN", None)
        del Cb
        for variant in ((0, 1) if not a.quick else (0,)) if a.policies == 2 else ():
            pkg.set_tuning("CSR_VARIANT", variant)
            for budget_mb in (48, 64, 80, 96) if not a.quick else (64,):
                rows = budget_mb * (1 << 20) // (K * 2)
                hot, n_hot = pkg.csr_hot_columns(n, ci, rows)
                for pol in (1, 0) if not a.quick else (1,):
                    pkg.set_tuning("CSR_COLD_POLICY", pol)
                    run(f"bf16 B variant {variant} hot {budget_mb} MB cold_policy {pol}",
                        lambda: pkg.csrmm_ex(rp, ci, val, Bb, Cm, n, n, K, hot=hot), alg_b, n_hot=n_hot)
                    pkg.set_tuning("CSR_COLD_POLICY", None)
            pkg.set_tuning("CSR_VARIANT", None)
        del rp, ci, B, Bb, val, Cm, Cref
        torch.cuda.empty_cache()
    out.close()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py -- the driver-facing benchmark (contract: DESIGN.md section 6).

    python bench.py --gpus N --steps K --warmup W            # our engine
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU SpMM (spmm.cc csr_spmm)

A "step" is one pass of the hot path, C = A * B, over one synthetic graph.  The default workload is BASELINE.json
configs[4]: the ogbn-products-shaped power-law graph (2,449,029 nodes, 61,859,140 edges, columns drawn
preferentially, K = 256, fp32, CSR), sharded by nnz-balanced row ranges over the N GPUs.  The other configs run with
--workload / --op (one GPU):

    --workload c1_uniform | c2_arxiv | c3_proteins | c3_proteins_shuffled | c3_proteins_rcm | c3_proteins_rcm_gpu |
               c3_proteins_gorder |
               c4_molhiv | c5_products | c5_products_local
    --op       csr | csr_bf16b | bsr16 | bsr32 | hybrid16 | hybrid32 | block_gemm16 | block_gemm32   [--math tf32|bf16|fp32]

Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name -> (graphs config, generator overrides, K, description)
    "c1_uniform": ("c1_uniform", {}, 64, "configs[0]: random graph, 65,536 nodes, 16 distinct uniform columns per row"),
    "c2_arxiv": ("c2_arxiv", {}, 128, "configs[1]: power-law graph of ogbn-arxiv shape (169,343 nodes, 1,166,243 edges)"),
    "c3_proteins": ("c3_proteins", {}, 128,
                    "configs[2]: planted-community graph of ogbn-proteins shape (132,534 nodes, 39,561,252 edges) in "
                    "its planted (ideal) order"),
    "c3_proteins_shuffled": ("c3_proteins", dict(shuffle=True), 128, "configs[2] with node ids shuffled, no reordering"),
    "c3_proteins_rcm": ("c3_proteins", dict(shuffle=True), 128,
                        "configs[2] shuffled, then reordered by the reference's reverseCuthillMcKee "
                        "(reorder_strategy.cc:57-98; permutation from tests/golden/c3_orders.npz)"),
    "c3_proteins_rcm_gpu": ("c3_proteins", dict(shuffle=True), 128,
                            "configs[2] shuffled, then reordered ON THE DEVICE by spmmdb_order_rcm (the reference's "
                            "reverseCuthillMcKee, reorder_strategy.cc:57-98, with equal-degree neighbours in ascending id)"),
    "c3_proteins_gorder": ("c3_proteins", dict(shuffle=True), 128,
                           "configs[2] shuffled, then reordered by the reference's vendored Gorder (Gorder/main.cpp, "
                           "window 5; permutation from tests/golden/c3_orders.npz)"),
    "c4_molhiv": ("c4_molhiv", {}, 300, "configs[3]: block-diagonal batch of 41,127 molecule-sized graphs"),
    "c5_products": ("c5_products", {}, 256,
                    "configs[4]: power-law graph of ogbn-products shape (2,449,029 nodes, 61,859,140 edges), columns "
                    "drawn preferentially (SURVEY 8d)"),
    "c5_products_local": ("c5_products", dict(p_local=0.9, window=4096), 256,
                          "configs[4] shape with 90 % of a row's columns within ~4096 of the row (the locality a "
                          "community-aware reordering leaves behind)"),
}
OPS = ["csr", "csr_bf16b", "bsr16", "bsr32", "hybrid16", "hybrid32", "block_gemm16", "block_gemm32"]
MATH = {"fp32": 0, "tf32": 1, "bf16": 2}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return {}, 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi style clock / throttle sampling during the timed region (NVML)."""

    def __init__(self, index):
        self.samples, self.reasons, self.stop = [], set(), False
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
                 "sw_thermal_slowdown": 0x20, "hw_power_brake": 0x80}
        while not self.stop and nv is not None:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.005)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=1)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def reference_cpu_run(rp, ci, B, K, steps, warmup, budget_s):
    """Time the reference's own csr_spmm (spmm.cc:7-25 via oracle/_ref) -- or the oracle's restatement of it when
    oracle/_ref is absent -- on a bounded row sample of the workload.  Index widening (the reference takes int64)
    and the output buffer are made once, outside the timed calls, like spmm.cc:63-72 times the call only."""
    import numpy as np
    from oracle import oracle as O
    O.build()
    use_ref = O.ref_spmm_lib() is not None
    if use_ref:  # all host cores, also under torchrun (which exports OMP_NUM_THREADS=1 to its workers)
        O.ref_set_omp_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    cores = O.ref_omp_threads() if use_ref else (os.cpu_count() or 1)
    n = rp.size - 1
    dense = np.ascontiguousarray(B, np.float64)
    rp64 = rp.astype(np.int64)

    def prepare(rows):
        e = int(rp64[rows])
        return rp64[:rows + 1].copy(), ci[:e].astype(np.int64), np.zeros((rows, K), np.float64), e

    def run(prep):
        ip, ix, out, _ = prep
        t = time.perf_counter()
        if use_ref:
            O.ref_csr_spmm(ip, ix, dense, out=out)
        else:
            O.spmm_cc_csr(ip, ix, dense, out=out)
        return time.perf_counter() - t

    r0 = min(n, 4096)
    p0 = prepare(r0)
    t0 = run(p0)
    per_step = budget_s / max(1, steps + warmup)
    rows = int(min(n, max(r0, r0 * per_step / max(t0, 1e-6))))
    prep = prepare(rows)
    ts = []
    for i in range(warmup + steps):
        t = run(prep)
        if i >= warmup:
            ts.append(t)
    tavg = sum(ts) / len(ts) if ts else t0
    e = prep[3]
    gflops = 2.0 * e * K / tavg / 1e9
    whole = "the whole workload" if rows == n else f"first {rows} rows ({e} nnz) of the workload"
    return dict(value=gflops, unit="GFLOP/s", cores=cores, kind="reference" if use_ref else "port",
                sample=f"{whole}, K={K}, fp64 unit weights (spmm.cc csr_spmm), {len(ts)} timed passes"), tavg * 1e3


def vendor_numbers(workload, op):
    """Same-box-class comparator timings recorded by baseline/compare.py (profiles/r2_vendor.jsonl): the cuSPARSE /
    cuBLAS calls the reference makes and the reference's own kernels compiled for sm_100a.  Not measured in this run."""
    p = os.path.join(ROOT, "profiles", "r2_vendor.jsonl")
    if not os.path.exists(p):
        return None
    key = workload.split("_")[0]
    want_op = "csrmm" if op.startswith("csr") else "bsrmm"
    bs = int("".join(ch for ch in op if ch.isdigit()) or 0) if not op.startswith("csr") else 0
    out = {}
    for line in open(p):
        try:
            d = json.loads(line)
        except ValueError:
            continue
        if d.get("config") == key and d.get("op") == want_op and "ms_med" in d and (not bs or d.get("bs") == bs):
            out[d["impl"]] = round(d["ms_med"], 4)
    return {"source": "profiles/r2_vendor.jsonl (baseline/compare.py, B200, not this run)", "ms": out} if out else None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c5_products", choices=sorted(WORKLOADS))
    ap.add_argument("--op", default="csr", choices=OPS)
    ap.add_argument("--math", default="tf32", choices=sorted(MATH))
    ap.add_argument("--density", type=float, default=0.1, help="hybrid ops: block fill threshold of the split")
    ap.add_argument("--scale", type=float, default=1.0, help="shrink the graph (debug only; invalidates the number)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-budget", type=float, default=20.0, help="seconds of CPU work for the cpu baseline")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "ours" else a.warmup

    import numpy as np
    import torch
    from spmm_denseblock_b200 import graphs

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cfg_name, overrides, K, desc = WORKLOADS[a.workload]
    reordered = a.workload in ("c3_proteins_rcm", "c3_proteins_gorder", "c3_proteins_rcm_gpu")
    cpu_graph = "shuffle" in overrides  # the permutation fixtures belong to the CPU generator's graph
    big = a.workload.startswith(("c5", "c4"))
    config = {"workload": f"{a.workload}: {desc}; op {a.op}", "K": K, "op": a.op,
              "generator": dict(overrides, seed=1234, scale=a.scale, device="cpu" if cpu_graph else "cuda"),
              "parallelism": f"row-shard x{world} (nnz-balanced), B replicated by all-gather",
              "l2": "inputs_exceed_l2 (B + C + A exceed the 126 MB L2)" if big
                    else "l2_flushed (256 MB written between timed steps)"}
    flush_l2 = not big

    def make_graph(device):
        return graphs.make_config_graph(cfg_name, device="cpu" if cpu_graph else device, scale=a.scale, **overrides)

    if a.impl == "reference":
        if rank != 0:
            return
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        rp, ci = make_graph(dev)
        n = rp.numel() - 1
        g = torch.Generator().manual_seed(1234)
        B = (torch.rand(n, K, generator=g) * 2 - 1).numpy()
        cb, ms = reference_cpu_run(rp.cpu().numpy(), ci.cpu().numpy(), B, K, a.steps, a.warmup,
                                   budget_s=min(150.0, 8.0 * (a.steps + a.warmup)))
        print(json.dumps({
            "impl": "reference", "metric": "SpMM GFLOP/s (2*nnz*K/t)", "value": cb["value"], "unit": "GFLOP/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}))
        return

    import spmm_denseblock_b200 as pkg
    from spmm_denseblock_b200 import shard
    L = pkg.lib()  # raises if the CUDA library is missing: there is no fallback
    assert torch.cuda.is_available(), "bench.py (impl ours) needs a CUDA device"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        assert a.op == "csr", "multi-GPU runs shard the CSR path (BASELINE.json configs[4])"
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    stream = torch.cuda.current_stream()
    sp = C.c_void_p(stream.cuda_stream)
    pk, peak, peak_src = peaks()
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if flush_l2 else None
    extra = {}

    # ---- workload: identical on every rank (same seed) ------------------------------------------------------------
    rp, ci = make_graph(dev)
    rp, ci = rp.to(dev), ci.to(dev)
    n, nnz = rp.numel() - 1, ci.numel()
    val = torch.ones(nnz, device=dev)  # unit adjacency weights (run_csrmm.cu:66, run_bsrmm.cu:66)
    g = torch.Generator(device=dev).manual_seed(1234)
    B = torch.rand(n, K, generator=g, device=dev) * 2 - 1  # U(-1,1) (load_data.cc:29-40)
    graph_info = {"rows": n, "nnz": nnz, "max_degree": int((rp[1:] - rp[:-1]).max().item())}

    # ---- config 3 as BASELINE.json states it: CPU reorderer -> permutation -> (csr2bsr ->) SpMM -------------------
    unperm = None
    if reordered:
        if a.workload.endswith("_gpu"):
            pkg.order_rcm(rp, ci)  # warm-up (scratch, kernels)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            old2new = pkg.order_rcm(rp, ci)  # C ABI: spmmdb_order_rcm (synchronises)
            extra["order_rcm_ms_once"] = (time.perf_counter() - t0) * 1e3
        else:
            fx = np.load(os.path.join(ROOT, "tests", "golden", "c3_orders.npz"))
            fp = np.array([n, nnz, int(rp.cpu().numpy()[::997].astype(np.int64).sum() % (1 << 31)),
                           int((ci.cpu().numpy()[::9973].astype(np.int64) * 31 % 1000003).sum() % (1 << 31))], np.int64)
            assert a.scale == 1.0 and np.array_equal(fp, fx["fingerprint"]), "c3_orders.npz belongs to another graph"
            old2new = torch.from_numpy(fx["old2new_" + a.workload.rsplit("_", 1)[1]].astype(np.int32)).to(dev)
        rp0, ci0, B0 = rp, ci, B
        t0 = time.perf_counter()
        rp, ci, val = pkg.permute_csr(rp0, ci0, val, old2new)   # A' = P A P^T on the device
        B = pkg.permute_rows(B0, old2new)                       # B' = P B
        torch.cuda.synchronize()
        extra["permute_ms_once"] = (time.perf_counter() - t0) * 1e3
        new2old = torch.empty_like(old2new)
        new2old[old2new.long()] = torch.arange(n, device=dev, dtype=torch.int32)
        unperm = (rp0, ci0, B0, new2old)

    # ---- the op --------------------------------------------------------------------------------------------------
    bs = int("".join(ch for ch in a.op if ch.isdigit()) or 0) if not a.op.startswith("csr") else 0
    math = MATH[a.math]
    launches_per_step = 2
    hyb = None
    if a.op in ("csr", "csr_bf16b"):
        ranges = shard.shard_plan(rp, world)  # C ABI: spmmdb_shard_plan
        lo, hi = ranges[rank]
        rps, cis, vals = shard.shard_csr(rp, ci, val, lo, hi)
        ms_, nnzs = hi - lo, cis.numel()
        Cs = torch.empty(ms_, K, device=dev)
        # every rank cuts long rows like the whole matrix would be cut: shards concatenate bit for bit
        seg = pkg.csrmm_segment(nnz, m=n)
        pkg.set_csrmm_segment(seg)
        graph_info.update(shard_rows=ms_, shard_nnz=nnzs, segment=seg)
        if a.op == "csr":
            def step():
                rc = L.spmmdb_csrmm_f32(ms_, n, K, nnzs, rps.data_ptr(), cis.data_ptr(), vals.data_ptr(), B.data_ptr(),
                                        K, 0, Cs.data_ptr(), K, 0, 1.0, 0.0, sp)
                assert rc == 0, rc
            alg_bytes = 4 * (ms_ + 1) + 8 * nnzs + 4 * n * K + 4 * ms_ * K  # B read once, C written once
            gather = 4 * nnzs * K
            dtype = "f32"
            small = nnzs < (4 << 20)
            kernel_name = ("csrmm_rows_pipe_kernel + long-row kernels (small-matrix path)" if small and 32 < K <= 256 else
                           "csrmm_rows_kernel" if nnzs // max(ms_, 1) >= 64 else
                           "csrmm_lean_kernel (8-aligned gather batches, queued long-row segments taken between items)"
                           if K <= 256 else "csrmm_stream_kernel (queued long-row segments taken between items)")
            launches_per_step = 5 if small and 32 < K <= 256 else 2
        else:
            t0 = time.perf_counter()
            Bb = pkg.dense_pack_bf16(B)
            torch.cuda.synchronize()
            extra["pack_B_bf16_ms_once"] = (time.perf_counter() - t0) * 1e3

            def step():
                rc = L.spmmdb_csrmm_ex_f32(ms_, n, K, nnzs, rps.data_ptr(), cis.data_ptr(), vals.data_ptr(),
                                           Bb.data_ptr(), 1, K, Cs.data_ptr(), K, 1.0, 0.0, None, sp)
                assert rc == 0, rc
            alg_bytes = 4 * (ms_ + 1) + 8 * nnzs + 2 * n * K + 4 * ms_ * K
            gather = 2 * nnzs * K
            dtype = "f32 accumulate, bf16-stored B"
            kernel_name = "csrmm_lean_kernel (bf16-stored B)" if K <= 512 else "csrmm_stream_kernel (bf16-stored B)"
        flops = 2.0 * nnz * K
        out_rows = Cs
        roof_extra = {"gather_bytes_per_launch": gather,
                      "note": "B-row gathers (nnz*K*e B) are served by L2/L1, not counted as algorithmic bytes"}
    else:
        assert world == 1
        mb = (n + bs - 1) // bs
        mp = mb * bs
        Bp = torch.zeros(mp, K, device=dev)
        Bp[:n] = B
        Cs = torch.zeros(mp, K, device=dev)
        t0 = time.perf_counter()
        hyb = None
        if a.op.startswith("hybrid"):
            hyb = pkg.Hybrid(n, n, rp, ci, val, bs, a.density)  # spmmdb_hybrid_create: the split, kept on the device
            nnzb, rem = hyb.nnzb, hyb.csr_nnz
        else:
            brp, bci, bv = pkg.csr2bsr(n, n, rp, ci, val, bs)
            nnzb, rem = bci.numel(), 0
        torch.cuda.synchronize()
        extra["conversion_ms_once"] = (time.perf_counter() - t0) * 1e3
        in_blocks = nnz - rem
        graph_info.update(block_size=bs, nnzb=nnzb, fill=in_blocks / max(nnzb * bs * bs, 1),
                          nnz_in_blocks=in_blocks, remainder_nnz=rem)
        if a.op.startswith("block_gemm"):
            bvc = bv.view(nnzb, bs, bs).transpose(1, 2).contiguous().view(-1)  # column-major blocks (block_cublas.cu)
            Zc = torch.zeros(K, mp, device=dev)

            def step():
                Zc.zero_()  # alpha = beta = 1 accumulate into a zeroed z (block_cublas.cu:118-136)
                rc = L.spmmdb_block_gemm_f32(mb, K, nnzb, bs, brp.data_ptr(), bci.data_ptr(), bvc.data_ptr(),
                                             Bp.data_ptr(), Zc.data_ptr(), math, sp)
                assert rc == 0, rc
            out_rows = Zc
            kernel_name = "bsrmm_tc_kernel (grouped block GEMM)" if math else "bsrmm_simt_kernel"
            launches_per_step = 4
        elif a.op.startswith("hybrid"):
            Cs = torch.zeros(n, K, device=dev)

            def step():
                # divide.cu:348-373 behind one call: C = A_csr * B, then C += A_bsr * B (spmmdb_hybrid_spmm_f32);
                # B and C keep their n rows, no padding to whole blocks
                hyb.spmm(B, Cs, K, math=math)
            out_rows = Cs
            kernel_name = "csrmm (remainder) + bsrmm_tc_kernel (dense blocks) through spmmdb_hybrid_spmm_f32"
            launches_per_step = 6
        else:
            def step():
                rc = L.spmmdb_bsrmm_f32(0, 1, mb, K, mb, nnzb, bs, 1.0, bv.data_ptr(), brp.data_ptr(), bci.data_ptr(),
                                        Bp.data_ptr(), K, 0.0, Cs.data_ptr(), K, 0, math, sp)
                assert rc == 0, rc
            out_rows = Cs
            kernel_name = "bsrmm_tc_kernel" if math else "bsrmm_simt_kernel"
            launches_per_step = 3
        flops = 2.0 * nnz * K
        dense_flops = 2.0 * nnzb * bs * bs * K
        alg_bytes = 4 * (mb + 1) + 4 * nnzb + 4 * nnzb * bs * bs + 4 * mp * K + 4 * mp * K
        if hyb is not None:
            alg_bytes += 4 * (n + 1) + 8 * rem
        dtype = {0: "f32", 1: "tf32 inputs, f32 accumulate", 2: "bf16 inputs, f32 accumulate"}[math]
        roof_extra = {"dense_flops_per_launch": dense_flops}
        ms_, nnzs, lo, hi = n, nnz, 0, n

    for _ in range(a.warmup):
        step()
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    with ClockSampler(local) as clk:
        for i in range(a.steps):
            if flush_buf is not None:
                flush_buf.zero_()
            ev[i][0].record()
            step()
            ev[i][1].record()
        barrier()
    step_ms = [s.elapsed_time(e) for s, e in ev]
    total_ms = sum(step_ms)
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / a.steps
    value = flops / (ms_per_step * 1e-3) / 1e9
    checksum = float(out_rows.double().sum().item())

    # ---- roofline of the dominant kernel (this rank's launch) ------------------------------------------------------
    avg_ms = total_ms / len(step_ms)
    achieved = alg_bytes / (avg_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "csrmm_traffic.json")
    if os.path.exists(tp) and world == 1 and a.scale == 1.0:
        traffic = json.load(open(tp)).get(f"{a.workload}:{a.op}")
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": kernel_name, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes}
    roofline.update(roof_extra)
    if traffic:
        roofline["dram_traffic_rate_gbs"] = traffic / (avg_ms * 1e-3) / 1e9
        roofline["dram_traffic_frac_of_peak"] = roofline["dram_traffic_rate_gbs"] / peak
    if bs:
        tens = None
        tpk = os.path.join(ROOT, "profiles", "r2_tf32_peak.json")
        if os.path.exists(tpk):
            d = json.load(open(tpk))
            tens = d.get("tf32_tflops") if math == 1 else d.get("bf16_tflops")
        if tens is None and math == 2:
            tens = pk.get("bf16_tflops")
        dt = dense_flops / (avg_ms * 1e-3) / 1e12
        roofline["tensor"] = {"dense_tflops": dt, "useful_tflops": flops / (avg_ms * 1e-3) / 1e12, "peak_tflops": tens,
                              "frac_of_tensor_peak": (dt / tens) if tens else None,
                              "note": "at K = 128 the stream of fp32-stored blocks caps dense throughput at 2K/4 "
                                      "flop/B x HBM = ~419 TFLOP/s: the op is HBM-bound on A, not tensor-bound"}

    # ---- result check outside the timed region ---------------------------------------------------------------------
    if unperm is not None:
        rp0, ci0, B0, new2old = unperm
        Cref = torch.empty(n, K, device=dev)
        pkg.set_csrmm_segment(0)
        pkg.csrmm(rp0, ci0, torch.ones(nnz, device=dev), B0, Cref, n, n, K)
        got = out_rows.t()[:n].contiguous() if a.op.startswith("block_gemm") else out_rows[:n].contiguous()
        Cun = pkg.permute_rows(got, new2old)  # C = P^T C'
        scale = float(Cref.abs().max().item())
        extra["unpermuted_result_max_abs_err_over_max_abs"] = float((Cun - Cref).abs().max().item()) / max(scale, 1e-30)
        del Cref, Cun

    # ---- all-gather of B and the fused all-gather + SpMM (C entry, NCCL) -------------------------------------------
    mg = None
    if world > 1:
        mg = shard.MultiGpu(rank, world)
        Bown = B[lo:hi].contiguous()
        Bg = torch.empty(n, K, device=dev)

        def timed(fn, reps=5):
            best = None
            for _ in range(reps):
                barrier()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                fn()
                e.record()
                barrier()
                tt = torch.tensor([s.elapsed_time(e)], device=dev, dtype=torch.float64)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                best = float(tt.item()) if best is None else min(best, float(tt.item()))
            return best

        def gather_only():
            Bg[lo:hi].copy_(Bown)
            mg.allgather_rows(Bg, ranges)
        extra["allgather_B_ms"] = timed(gather_only)
        Cf = torch.empty(ms_, K, device=dev)
        for slabs in (1, 2, 4):
            extra[f"allgather_plus_spmm_ms_slabs{slabs}"] = timed(
                lambda: mg.sharded_csrmm(rps, cis, vals, Bown, ranges, Cf, n, K, slabs=slabs))
        extra["sharded_entry_matches_replicated"] = bool(torch.equal(Cf, Cs))
        extra["allgather_path"] = {0: "NCCL send / recv group", 1: "CUDA IPC peer pushes by copy engines + NCCL barriers",
                                   2: "pack kernel storing into every rank's replica over NVLink (CUDA IPC) + NCCL "
                                      "barriers"}[pkg.get_tuning("MG_PEER", 2)]
        best = min(extra[f"allgather_plus_spmm_ms_slabs{s_}"] for s_ in (1, 2, 4))
        extra["value_with_allgather"] = flops / (best * 1e-3) / 1e9
        extra["allgather_inbound_gbs"] = 4.0 * (n - ms_) * K / (extra["allgather_B_ms"] * 1e-3) / 1e9

    # ---- e2e: HOST buffers (pinned) in, HOST buffer out, copies inside the timed region ------------------------------
    e2e = None
    if a.e2e_steps > 0:
        hC = torch.empty(out_rows.shape).pin_memory()
        out_rows.zero_()  # the e2e result must come from the e2e steps, not from the device-timed run above
        if a.op == "csr" and world == 1:
            h = {k_: v_.cpu().pin_memory() for k_, v_ in dict(rp=rps, ci=cis, val=vals, B=B).items()}
            api = "spmmdb_csrmm_host_f32 (pinned host buffers; feature slabs pipelined over H2D / kernel / D2H)"
            h2d = 4 * (ms_ + 1) + 8 * nnzs + 4 * n * K

            def e2e_step():
                rc = L.spmmdb_csrmm_host_f32(ms_, n, K, nnzs, h["rp"].data_ptr(), h["ci"].data_ptr(),
                                             h["val"].data_ptr(), h["B"].data_ptr(), hC.data_ptr(), 1.0, 0.0, sp)
                assert rc == 0, rc
        elif a.op == "csr":
            # every rank owns its rows of B (the C rows it produced as the previous layer's output):
            # H2D of its A shard and its B rows, spmmdb_sharded_csrmm_f32 (all-gather-v over NVLink overlapped with
            # the SpMM per feature slab), D2H of its C rows
            h = {k_: v_.cpu().pin_memory() for k_, v_ in dict(rp=rps, ci=cis, val=vals, B=B[lo:hi]).items()}
            d = {k_: torch.empty_like(v_, device=dev) for k_, v_ in h.items()}
            # Pipelined over feature slabs like the one-GPU host entry: the H2D copy of slab s + 1, the all-gather +
            # SpMM of slab s and the D2H copy of slab s - 1 run on three streams (PCIe is full duplex, and the ranks
            # share the host's PCIe switches and memory, so the copies are what the step costs).
            n_sl = 4 if K % 256 == 0 else (2 if K % 128 == 0 else 1)
            Ks = K // n_sl
            api = (f"row shards, {n_sl} feature slabs in a pipeline: H2D of this rank's A shard and of a slab of its B rows "
                   "(pinned, cudaMemcpy2DAsync), spmmdb_sharded_csrmm_f32 on the slab (all-gather-v of B over NVLink + SpMM), "
                   "D2H of the slab of its C rows")
            h2d = 4 * (ms_ + 1) + 8 * nnzs + 4 * ms_ * K
            rt = C.CDLL("libcudart.so.12")
            rt.cudaMemcpy2DAsync.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int,
                                             C.c_void_p]
            s_run, s_out = torch.cuda.Stream(), torch.cuda.Stream()
            ev_in = [torch.cuda.Event() for _ in range(n_sl)]
            ev_run = [torch.cuda.Event() for _ in range(n_sl)]
            marks = [torch.cuda.Event(enable_timing=True) for _ in range(4)]

            def copy2d(dst, src, rows, kind, st):
                rc = rt.cudaMemcpy2DAsync(dst, 4 * K, src, 4 * K, 4 * Ks, rows, kind, C.c_void_p(st.cuda_stream))
                assert rc == 0, rc

            def e2e_step():
                marks[0].record(stream)
                for k_ in ("rp", "ci", "val"):
                    d[k_].copy_(h[k_], non_blocking=True)
                for sl in range(n_sl):
                    off = 4 * sl * Ks
                    if ms_ > 0:
                        copy2d(d["B"].data_ptr() + off, h["B"].data_ptr() + off, ms_, 1, stream)
                    ev_in[sl].record(stream)
                    s_run.wait_event(ev_in[sl])
                    with torch.cuda.stream(s_run):
                        mg.sharded_csrmm(d["rp"], d["ci"], d["val"], d["B"][:, sl * Ks:], ranges, Cs[:, sl * Ks:], n, Ks,
                                         slabs=1)
                    ev_run[sl].record(s_run)
                    s_out.wait_event(ev_run[sl])
                    if ms_ > 0:
                        copy2d(hC.data_ptr() + off, Cs.data_ptr() + off, ms_, 2, s_out)
                marks[1].record(stream)
                marks[2].record(s_run)
                marks[3].record(s_out)
                stream.synchronize()
                s_run.synchronize()
                s_out.synchronize()
        else:
            # the op's device inputs from pinned host copies, the device call(s) through the C ABI, C back to the host
            if a.op == "csr_bf16b":
                names = dict(rp=rps, ci=cis, val=vals, B=Bb)
            elif hyb is not None:
                names = dict(B=B)  # the split operator stays on the device (made once per graph); B is the step's input
            else:
                names = dict(brp=brp, bci=bci, B=Bp, bv=(bvc if a.op.startswith("block_gemm") else bv))
            h = {k_: v_.cpu().pin_memory() for k_, v_ in names.items()}
            api = f"pinned host copies of the {a.op} operands -> H2D -> C-ABI device call(s) -> D2H of C"
            h2d = sum(v_.numel() * v_.element_size() for v_ in h.values())

            def e2e_step():
                for k_, v_ in names.items():
                    v_.copy_(h[k_], non_blocking=True)
                step()
                hC.copy_(out_rows, non_blocking=True)
                stream.synchronize()
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.e2e_steps):
            e2e_step()
        barrier()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / a.e2e_steps
        tt = torch.tensor([e2e_ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_ms = float(tt.item())
        e2e_ok = abs(float(hC.double().sum().item()) - checksum) <= 1e-6 * max(1.0, abs(checksum))
        if world > 1:
            # where the step goes on this rank (device timeline of the last step; the ranks share the host's memory and
            # PCIe switches, which is why the copies run far below one GPU's PCIe rate)
            extra["e2e_breakdown_ms_rank0"] = {"h2d_done_at": marks[0].elapsed_time(marks[1]),
                                               "allgather_plus_spmm_done_at": marks[0].elapsed_time(marks[2]),
                                               "d2h_done_at": marks[0].elapsed_time(marks[3]),
                                               "h2d_GBps": h2d / marks[0].elapsed_time(marks[1]) / 1e6,
                                               "note": "device timeline of the last step, ms after its start; the three "
                                                       "phases overlap"}
        e2e = {"value": flops / (e2e_ms * 1e-3) / 1e9, "unit": "GFLOP/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": hC.numel() * 4, "ms_per_step": e2e_ms, "matches_device_path": bool(e2e_ok),
               "api": api}
        del h, hC

    out = {"metric": "SpMM GFLOP/s (2*nnz*K/t)", "value": value, "unit": "GFLOP/s", "n_gpus": world,
           "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": dtype, "data": "synthetic", "config": config,
           "graph": graph_info, "roofline": roofline, "e2e": e2e, "gpu_launches": launches_per_step * a.steps,
           "clocks": clk.summary(), "checksum": checksum}
    out.update(extra)
    v = vendor_numbers(a.workload, a.op)
    if v:
        out["comparators"] = v

    # second workload beside the default: the same shape with locality (what a reordering leaves behind)
    if rank == 0 and world == 1 and a.workload == "c5_products" and a.op == "csr" and not a.no_extras and a.scale == 1.0:
        del rps, cis, vals, Cs, out_rows, val
        torch.cuda.empty_cache()
        rp2, ci2 = graphs.make_config_graph("c5_products", device=dev, **WORKLOADS["c5_products_local"][1])
        v2 = torch.ones(ci2.numel(), device=dev)
        C2 = torch.empty(n, K, device=dev)
        pkg.set_csrmm_segment(pkg.csrmm_segment(ci2.numel(), m=n))
        ts = []
        for i in range(3 + 10):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            pkg.csrmm(rp2, ci2, v2, B, C2, n, n, K)
            e.record()
            e.synchronize()
            if i >= 3:
                ts.append(s.elapsed_time(e))
        m2 = sum(ts) / len(ts)
        b2 = 4 * (n + 1) + 8 * ci2.numel() + 8 * n * K
        out["second_workload"] = {"workload": "c5_products_local", "nnz": ci2.numel(), "ms_per_step": m2,
                                  "value": 2.0 * ci2.numel() * K / (m2 * 1e-3) / 1e9,
                                  "roofline_frac": b2 / (m2 * 1e-3) / 1e9 / peak}
        del rp2, ci2, v2, C2

    if rank == 0 and world == 1 and not a.no_cpu:
        src = unperm if unperm is not None else (rp, ci, B)
        cb, _ = reference_cpu_run(src[0].cpu().numpy(), src[1].cpu().numpy(), src[2].cpu().numpy(), K, 1, 0, a.cpu_budget)
        out["cpu_baseline"] = cb
    if rank == 0:
        print(json.dumps(out), flush=True)
    if mg is not None:
        mg.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py -- the driver-facing benchmark (see the contract in DESIGN.md section 6).

    python bench.py --gpus N --steps K --warmup W            # our engine
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU SpMM

A "step" is one pass of the hot path -- C = A * B, CSR x dense, fp32 -- over the
products-shaped synthetic graph (BASELINE.json configs[4]: 2,449,029 nodes, ~61.9 M
edges, K = 256), sharded by nnz-balanced row ranges over the N GPUs with B replicated.
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
    # name -> (graphs config, overrides, K)
    "c5_products": ("c5_products", dict(), 256),  # power-law, columns drawn preferentially
    "c5_products_local": ("c5_products", dict(p_local=0.9, window=4096), 256),  # 90 % of the columns near the row
    "c2_arxiv": ("c2_arxiv", dict(), 128),
    "c4_molhiv": ("c4_molhiv", dict(), 300),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


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
            time.sleep(0.02)

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
    """Time the reference's own csr_spmm (spmm.cc:7-25 via oracle/_ref) -- or the oracle's
    restatement of it when oracle/_ref is absent -- on a bounded row sample of the workload."""
    import numpy as np
    from oracle import oracle as O
    O.build()
    use_ref = O.ref_spmm_lib() is not None
    fn = O.ref_csr_spmm if use_ref else O.spmm_cc_csr
    if use_ref:  # all host cores, also under torchrun (which exports OMP_NUM_THREADS=1 to its workers)
        O.ref_set_omp_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    cores = O.ref_omp_threads() if use_ref else (os.cpu_count() or 1)
    n = rp.size - 1
    dense = B.astype(np.float64)
    rp64 = rp.astype(np.int64)

    def run_rows(rows):
        e = int(rp64[rows])
        t = time.perf_counter()
        fn(rp64[:rows + 1], ci[:e].astype(np.int64), dense)
        return time.perf_counter() - t, e

    r0 = min(n, 4096)
    t0, e0 = run_rows(r0)
    per_step = budget_s / max(1, steps + warmup)
    rows = int(min(n, max(r0, r0 * per_step / max(t0, 1e-6))))
    ts = []
    for i in range(warmup + steps):
        t, e = run_rows(rows)
        if i >= warmup:
            ts.append(t)
    tmax = max(ts) if ts else t0
    tavg = sum(ts) / len(ts) if ts else t0
    gflops = 2.0 * e * K / tavg / 1e9
    return dict(value=gflops, unit="GFLOP/s", cores=cores, kind="reference" if use_ref else "port",
                sample=f"first {rows} rows ({e} nnz) of the workload, K={K}, fp64 unit weights (spmm.cc csr_spmm), "
                       f"{len(ts)} timed passes"), tavg * 1e3, tmax


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c5_products", choices=sorted(WORKLOADS))
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
    cfg_name, overrides, K = WORKLOADS[a.workload]
    config = {"workload": f"{a.workload}: synthetic power-law graph of ogbn-products shape, CSR x dense fp32",
              "K": K, "generator": dict(overrides, seed=1234, scale=a.scale),
              "parallelism": f"row-shard x{world} (nnz-balanced), B replicated",
              "l2": "inputs_exceed_l2 (B and C are 2.5 GB each vs 126 MB L2)"}

    if a.impl == "reference":
        if rank != 0:
            return
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        rp, ci = graphs.make_config_graph(cfg_name, device=dev, scale=a.scale, **overrides)
        n = rp.numel() - 1
        g = torch.Generator().manual_seed(1234)
        B = (torch.rand(n, K, generator=g) * 2 - 1).numpy()
        cb, ms, _ = reference_cpu_run(rp.cpu().numpy(), ci.cpu().numpy(), B, K, a.steps, a.warmup,
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
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- workload: identical on every rank (same seed), then sharded ----------------
    rp, ci = graphs.make_config_graph(cfg_name, device=dev, scale=a.scale, **overrides)
    n, nnz = rp.numel() - 1, ci.numel()
    val = torch.ones(nnz, device=dev)  # unit adjacency weights (run_csrmm.cu:66)
    g = torch.Generator(device=dev).manual_seed(1234)
    B = torch.rand(n, K, generator=g, device=dev) * 2 - 1  # U(-1,1) (load_data.cc:29-40)
    ranges = shard.shard_plan(rp, world)  # C ABI: spmmdb_shard_plan
    lo, hi = ranges[rank]
    rps, cis, vals = shard.shard_csr(rp, ci, val, lo, hi)
    ms_, nnzs = hi - lo, cis.numel()
    Cs = torch.empty(ms_, K, device=dev)
    del val
    stream = torch.cuda.current_stream()

    # every rank cuts long rows like the whole matrix would be cut: shards concatenate bit for bit
    seg = pkg.csrmm_segment(nnz, m=n)
    pkg.set_csrmm_segment(seg)

    def step():
        rc = L.spmmdb_csrmm_f32(ms_, n, K, nnzs, rps.data_ptr(), cis.data_ptr(), vals.data_ptr(), B.data_ptr(), K, 0,
                                Cs.data_ptr(), K, 0, 1.0, 0.0, C.c_void_p(stream.cuda_stream))
        assert rc == 0, rc

    for _ in range(a.warmup):
        step()
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.steps + 1)]
    with ClockSampler(local) as clk:
        ev[0].record()
        for i in range(a.steps):
            step()
            ev[i + 1].record()
        barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(a.steps)]
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms_max = float(t.item())
    ms_per_step = total_ms_max / a.steps
    value = 2.0 * nnz * K / (ms_per_step * 1e-3) / 1e9
    checksum = float(Cs.double().sum().item())

    # ---- roofline of the dominant kernel (this rank's launch) ------------------------
    peak, peak_src = peaks()
    alg_bytes = 4 * (ms_ + 1) + 8 * nnzs + 4 * n * K + 4 * ms_ * K  # B read once, C written once
    avg_ms = sum(step_ms) / len(step_ms)
    achieved = alg_bytes / (avg_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "csrmm_traffic.json")
    if os.path.exists(tp) and world == 1 and a.scale == 1.0:
        traffic = json.load(open(tp)).get(a.workload)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": "csrmm_stream_kernel (queued long-row segments taken between items)", "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes,
                "gather_bytes_per_launch": 4 * nnzs * K,
                "note": "B-row gathers (nnz*K*4 B) are served by L2/L1, not counted as algorithmic bytes"}

    # ---- all-gather of B (how B gets replicated when each rank owns its rows of B) ----
    allgather_ms = None
    if world > 1:
        for it in range(3):
            barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            shard.allgather_rows(B, ranges)
            e.record()
            barrier()
            tt = torch.tensor([s.elapsed_time(e)], device=dev, dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            allgather_ms = float(tt.item())

    # ---- e2e: HOST buffers (pinned) in, HOST buffer out, copies inside the timed region ------------
    # N = 1: the C-ABI host entry (feature slabs pipelined over H2D / kernel / D2H).
    # N > 1: every rank owns its rows of B (the C rows it produced as the previous layer's output):
    #        H2D of its A shard and its B rows, all-gather-v of B over NVLink (NCCL), the device call, D2H of
    #        its C rows.  Copying the whole of B to every GPU instead makes all ranks pull 2.5 GB through the
    #        host's memory and PCIe switches at once (N = 8: 139 ms against 72 ms at N = 1).
    hC = torch.empty(ms_, K).pin_memory()
    Cs.zero_()  # the e2e result must come from the e2e steps, not from the device-timed run above
    if world == 1:
        h = {k_: v_.cpu().pin_memory() for k_, v_ in dict(rp=rps, ci=cis, val=vals, B=B).items()}
        e2e_api = "spmmdb_csrmm_host_f32 (pinned host buffers; feature slabs pipelined over H2D / kernel / D2H)"
        h2d = 4 * (ms_ + 1) + 8 * nnzs + 4 * n * K

        def e2e_step():
            rc = L.spmmdb_csrmm_host_f32(ms_, n, K, nnzs, h["rp"].data_ptr(), h["ci"].data_ptr(),
                                         h["val"].data_ptr(), h["B"].data_ptr(), hC.data_ptr(), 1.0, 0.0,
                                         C.c_void_p(stream.cuda_stream))
            assert rc == 0, rc
    else:
        h = {k_: v_.cpu().pin_memory() for k_, v_ in dict(rp=rps, ci=cis, val=vals, B=B[lo:hi]).items()}
        d = {k_: torch.empty_like(v_, device=dev) for k_, v_ in h.items() if k_ != "B"}
        dB = torch.empty(n, K, device=dev)
        e2e_api = ("row shards: H2D of this rank's A shard and B rows (pinned), NCCL all-gather-v of B over NVLink "
                   "(shard.allgather_rows), spmmdb_csrmm_f32, D2H of this rank's C rows")
        h2d = 4 * (ms_ + 1) + 8 * nnzs + 4 * ms_ * K

        def e2e_step():
            dB[lo:hi].copy_(h["B"], non_blocking=True)
            for k_ in d:
                d[k_].copy_(h[k_], non_blocking=True)
            shard.allgather_rows(dB, ranges)
            rc = L.spmmdb_csrmm_f32(ms_, n, K, nnzs, d["rp"].data_ptr(), d["ci"].data_ptr(), d["val"].data_ptr(),
                                    dB.data_ptr(), K, 0, Cs.data_ptr(), K, 0, 1.0, 0.0,
                                    C.c_void_p(stream.cuda_stream))
            assert rc == 0, rc
            hC.copy_(Cs, non_blocking=True)
            stream.synchronize()
    if a.e2e_steps > 0:
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.e2e_steps):
        e2e_step()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / max(1, a.e2e_steps)
    tt = torch.tensor([e2e_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    e2e_ms = float(tt.item())
    e2e_ok = abs(float(hC.double().sum().item()) - checksum) <= 1e-6 * max(1.0, abs(checksum))
    d2h = 4 * ms_ * K
    e2e = {"value": 2.0 * nnz * K / (e2e_ms * 1e-3) / 1e9, "unit": "GFLOP/s", "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms, "matches_device_path": bool(e2e_ok),
           "api": e2e_api}

    out = {"metric": "SpMM GFLOP/s (2*nnz*K/t)", "value": value, "unit": "GFLOP/s", "n_gpus": world,
           "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
           "graph": {"rows": n, "nnz": nnz, "max_degree": int((rp[1:] - rp[:-1]).max().item()),
                     "shard_rows": ms_, "shard_nnz": nnzs, "segment": seg},
           "roofline": roofline, "e2e": e2e, "gpu_launches": 2 * a.steps, "clocks": clk.summary(),
           "checksum": checksum}
    if allgather_ms is not None:
        out["allgather_B_ms"] = allgather_ms
        out["value_with_allgather"] = 2.0 * nnz * K / ((ms_per_step + allgather_ms) * 1e-3) / 1e9

    if rank == 0 and world == 1 and not a.no_cpu:
        del h, hC
        cb, _, _ = reference_cpu_run(rp.cpu().numpy(), ci.cpu().numpy(), B.cpu().numpy(), K, 1, 0, a.cpu_budget)
        out["cpu_baseline"] = cb
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/bin/bash
# round-2 run W: the shared-memory staged csr2bsr fill -- conversion tests on every path, then csr2bsr timings on C3 / C4
# with the staged kernel (default) and with the zero-fill + scatter kernel (SPMMDB_CONVERT_STAGE=0)
mkdir -p gpurun_out
# (the staged kernel is not in the tree any more: this harness is kept for the record of how r2w was measured)
python - <<'P' | tee gpurun_out/r2w_csr2bsr_stage.jsonl
import json, os, sys
sys.path.insert(0, os.getcwd())
import torch
import spmm_denseblock_b200 as pkg
from spmm_denseblock_b200 import graphs
pkg.lib()
peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"] if os.path.exists("MEASURED_PEAKS.json") else 6650.0
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timeit(fn, iters=7):
    fn(); ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); e.synchronize(); ts.append(s.elapsed_time(e))
    ts.sort(); return ts[len(ts) // 2]
for cfg in ("c3_proteins", "c4_molhiv"):
    rp, ci = graphs.make_config_graph(cfg, device="cuda")
    n, nnz = rp.numel() - 1, ci.numel()
    val = torch.ones(nnz, device="cuda")
    for bs in (16, 32):
        mb = (n + bs - 1) // bs
        brp = torch.empty(mb + 1, dtype=torch.int32, device="cuda")
        nnzb = pkg.csr2bsr_nnz(n, n, rp, ci, bs, brp)
        if nnzb * bs * bs * 4 > 40e9: continue
        bci = torch.empty(nnzb, dtype=torch.int32, device="cuda")
        bv = torch.empty(nnzb * bs * bs, device="cuda")
        import ctypes as C
        L = pkg.lib(); vp = C.c_void_p
        def fill():
            rc = L.spmmdb_csr2bsr_f32(0, n, n, vp(rp.data_ptr()), vp(ci.data_ptr()), vp(val.data_ptr()), bs, vp(brp.data_ptr()),
                                      vp(bci.data_ptr()), vp(bv.data_ptr()), vp(torch.cuda.current_stream().cuda_stream))
            assert rc == 0
        alg = 4 * (n + 1) + 8 * nnz + 4 * (mb + 1) + 4 * nnzb + 4 * nnzb * bs * bs
        res = {}
        for stage in (1, 0):
            pkg.set_tuning("CONVERT_STAGE", stage)
            os.environ["SPMMDB_CONVERT_BR"] = "1"
            t = timeit(fill)
            res[stage] = (t, bv.clone())
        pkg.set_tuning("CONVERT_STAGE", None)
        print(json.dumps(dict(config=cfg, bs=bs, nnzb=nnzb, fill=nnz / (nnzb * bs * bs), fill_pass_staged_ms=res[1][0],
                              fill_pass_scatter_ms=res[0][0], staged_hbm_frac=alg / res[1][0] / 1e6 / peak,
                              scatter_hbm_frac=alg / res[0][0] / 1e6 / peak, bits_equal=bool(torch.equal(res[1][1], res[0][1])))), flush=True)
        del bv, bci, res
        torch.cuda.empty_cache()
P

"""Localise the hybrid mismatch: which of (divide, csrmm on the remainder, bsrmm fp32, bsrmm tf32) is off."""
import sys, os, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import spmm_denseblock_b200 as pkg
from spmm_denseblock_b200 import graphs

def main():
    for p_in, bs, dens in ((0.9, 32, 0.1), (0.6, 16, 0.1), (0.9, 16, 0.1)):
        rp, ci = graphs.make_config_graph("c3_proteins", device="cuda", p_in=p_in)
        n, nnz, K = rp.numel() - 1, ci.numel(), 128
        mb = (n + bs - 1) // bs
        val = torch.ones(nnz, device="cuda")
        Bp = torch.rand(mb * bs, K, device="cuda") * 2 - 1
        A = torch.sparse_csr_tensor(rp.long(), ci.long(), val, size=(n, mb * bs))
        ref = (A.to(torch.float64) @ Bp.to(torch.float64))
        crp, cci, cv, brp, bci, bv = pkg.divide(n, n, rp, ci, None, bs, dens)
        Ac = torch.sparse_csr_tensor(crp.long(), cci.long(), cv, size=(n, mb * bs))
        ref_c = Ac.to(torch.float64) @ Bp.to(torch.float64)
        # BSR part as dense-expanded reference via bsr2csr
        c2rp, c2ci, c2v = pkg.bsr2csr(mb, mb, brp, bci, bv, bs)
        Ab = torch.sparse_csr_tensor(c2rp.long(), c2ci.long(), c2v, size=(mb * bs, mb * bs))
        ref_b = (Ab.to(torch.float64) @ Bp.to(torch.float64))[:n]
        sc = ref.abs().max().item()
        out = dict(p_in=p_in, bs=bs, dens=dens, nnzb=bci.numel(), csr_nnz=cci.numel(),
                   split_sum=float(((ref_c + ref_b) - ref).abs().max() / sc))
        C = torch.zeros(mb * bs, K, device="cuda")
        pkg.csrmm(crp, cci, cv, Bp, C, n, n, K, nnz=cci.numel())
        out["csr_part"] = float((C[:n].double() - ref_c).abs().max() / sc)
        for math in (0, 1, 2):
            for beta in (0.0, 1.0):
                Cb = torch.full((mb * bs, K), 1.0, device="cuda")
                pkg.bsrmm(0, 1, mb, K, mb, bci.numel(), 1.0, bv, brp, bci, bs, Bp, K, beta, Cb, K, c_layout=0, math=math)
                d = (Cb[:n].double() - (ref_b + beta)).abs()
                out[f"bsr_m{math}_b{int(beta)}"] = float(d.max() / sc)
                if d.max() / sc > 0.05:
                    bad_rows = torch.nonzero(d.max(1).values / sc > 0.05).squeeze(1)
                    br = torch.unique(bad_rows // bs)
                    deg = (brp[1:] - brp[:-1])
                    out[f"bad_m{math}_b{int(beta)}"] = dict(n_bad_rows=int(bad_rows.numel()), first_block_rows=br[:12].tolist(),
                                                            their_block_counts=deg[br[:12]].tolist(),
                                                            neighbours=deg[max(int(br[0]) - 4, 0):int(br[0]) + 8].tolist())
        print(json.dumps(out), flush=True)

main()

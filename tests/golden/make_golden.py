"""Generate tests/golden/*.npz by RUNNING THE REFERENCE'S OWN CPU CODE (oracle/_ref,
built from /root/reference by oracle/Makefile) on seeded inputs.  Run in the
build container (needs /root/reference); the .npz files are committed and are
what travels to the GPU box.

  python tests/golden/make_golden.py
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def random_csr(rng, m, n, avg_deg, hub=None):
    rows = []
    for i in range(m):
        d = rng.poisson(avg_deg)
        if hub is not None and i == hub[0]:
            d = hub[1]
        rows.append(np.sort(rng.choice(n, size=min(d, n), replace=False)))
    rowptr = np.zeros(m + 1, np.int32)
    rowptr[1:] = np.cumsum([len(r) for r in rows])
    colind = np.concatenate(rows).astype(np.int32) if rowptr[-1] else np.zeros(0, np.int32)
    return rowptr, colind


def main():
    O.build()
    assert O.ref_spmm_lib() is not None and O.ref_host_lib() is not None, "oracle/_ref not built"
    rng = np.random.default_rng(1234)

    # 1. the reference's csr_spmm (spmm.cc:7-25): unit weights, double
    rp, ci = random_csr(rng, 300, 257, 9, hub=(17, 200))
    dense = rng.uniform(-1, 1, size=(257, 48))
    out = O.ref_csr_spmm(rp, ci, dense)
    np.savez_compressed(os.path.join(HERE, "ref_csr_spmm.npz"), rowptr=rp, colind=ci, dense=dense, out=out)

    # 2. calculateNnzb (utility.cc:47-69) for the block sizes reorder_graph.cc:12-24 sweeps
    rp, ci = random_csr(rng, 1000, 1000, 6, hub=(3, 400))
    bss = np.array([2, 3, 4, 8, 16, 32, 64], np.int32)
    nnzb = np.array([O.ref_calculate_nnzb(rp, ci, int(b)) for b in bss], np.int64)
    np.savez_compressed(os.path.join(HERE, "ref_calculate_nnzb.npz"), rowptr=rp, colind=ci, bs=bss, nnzb=nnzb)

    # 3. permutate (reorder_strategy.cc:26-39) and the RCM-like reordering (:57-98)
    rp, ci = random_csr(rng, 500, 500, 5)
    # make it symmetric like the graph dumps
    import scipy.sparse as sp
    A = sp.csr_matrix((np.ones(ci.size), ci, rp), shape=(500, 500))
    A = ((A + A.T) > 0).astype(np.float32).tocsr()
    A.sort_indices()
    rp, ci = A.indptr.astype(np.int32), A.indices.astype(np.int32)
    perm = rng.permutation(500).astype(np.int32)
    prp, pci = O.ref_permutate(rp, ci, perm)
    rrp, rci = O.ref_rcmk_graph(rp, ci)
    np.savez_compressed(os.path.join(HERE, "ref_permutate.npz"), rowptr=rp, colind=ci, old2new=perm,
                        p_rowptr=prp, p_colind=pci, rcmk_rowptr=rrp, rcmk_colind=rci)

    # 4. text formats (load_data.cc:125-184): dump with the reference, keep the bytes
    L = O.ref_host_lib()
    import ctypes as C
    with tempfile.TemporaryDirectory() as d:
        prefix = os.path.join(d, "toy")
        rp4, ci4 = random_csr(rng, 12, 12, 3)
        L.ref_dump_csr.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.ref_dump_csr(prefix.encode(), 12, int(ci4.size), rp4.ctypes.data, ci4.ctypes.data)
        indptr_txt = open(prefix + "_indptr.txt").read()
        indices_txt = open(prefix + "_indices.txt").read()
        # the toy edge list through loadGraphFromFile + convertGraphToCSR
        edge_txt = "6 10\n0 1\n1 0\n1 2\n2 1\n2 3\n3 2\n4 5\n5 4\n0 5\n5 0\n"
        open(os.path.join(d, "g.txt"), "w").write(edge_txt)
        grp = np.zeros(16, np.int32)
        gci = np.zeros(32, np.int32)
        nnz = C.c_int(0)
        L.ref_load_graph_csr.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        n = L.ref_load_graph_csr(os.path.join(d, "g.txt").encode(), grp.ctypes.data, 16, gci.ctypes.data, 32,
                                 C.byref(nnz))
    np.savez_compressed(os.path.join(HERE, "ref_text_formats.npz"), rowptr=rp4, colind=ci4,
                        indptr_txt=np.array(indptr_txt), indices_txt=np.array(indices_txt),
                        edge_txt=np.array(edge_txt), g_rowptr=grp[:n + 1], g_colind=gci[:nnz.value])

    # 5. divide_matrix (divide.cu:52-127): hybrid dense-block split on a graph with planted dense
    #    diagonal blocks (ragged last block row: 205 rows, bs 8 / 16 / 3), several densities
    assert O.ref_divide_lib() is not None, "oracle/_ref/libref_divide.so not built"
    n5 = 205
    rows = []
    for i in range(n5):
        c = set(rng.choice(n5, size=int(rng.poisson(4)), replace=False).tolist())
        if (i // 16) % 2 == 0:  # every other 16-row band has a dense diagonal block
            lo = (i // 16) * 16
            c |= set(j for j in range(lo, min(lo + 16, n5)) if rng.uniform() < 0.7)
        rows.append(np.array(sorted(c), np.int32))
    rp5 = np.zeros(n5 + 1, np.int32)
    rp5[1:] = np.cumsum([len(r) for r in rows])
    ci5 = np.concatenate(rows).astype(np.int32)
    out = dict(rowptr=rp5, colind=ci5)
    cases = [(8, 0.5), (16, 0.25), (16, 0.6), (3, 0.34), (16, 1.0), (32, 0.05)]
    out["cases"] = np.array(cases, np.float64)
    for k, (bs, dens) in enumerate(cases):
        crp, cci, brp, bci, bv = O.ref_divide_matrix(rp5, ci5, bs, dens)
        out.update({f"c{k}_csr_rowptr": crp, f"c{k}_csr_colind": cci, f"c{k}_bsr_rowptr": brp,
                    f"c{k}_bsr_colind": bci, f"c{k}_bsr_val": bv})
    np.savez_compressed(os.path.join(HERE, "ref_divide.npz"), **out)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()

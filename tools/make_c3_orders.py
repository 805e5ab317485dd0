#!/usr/bin/env python
"""Offline input preparation for BASELINE.json config 3 ("Rabbit/Gorder-reordered csr2bsr + BSR SpMM"), run in
the build container where /root/reference exists:

  planted-community proteins-shaped graph with SHUFFLED node ids (graphs.make_config_graph("c3_proteins",
  shuffle=True), CPU generator, seed 1234)
    -> the reference's own reverseCuthillMcKee (reorder_strategy.cc:57-98, through oracle/_ref/libref_host.so)
    -> the reference's vendored Gorder binary (/root/reference/Gorder/Gorder, main.cpp:25-75, window 5)
  Both emit the RELABELLED GRAPH, not the permutation (reorder_strategy.cc:26-39, Gorder/Graph.cpp:228-263), so the
  permutation old -> new is recovered here by colour refinement (isomorphism-invariant vertex signatures) and
  verified exactly: permuting the input must reproduce the emitted graph edge for edge.

Writes tests/golden/c3_orders.npz (old2new per ordering + a fingerprint of the graph they belong to).  The GPU box
regenerates the same graph (CPU generator, same seed), checks the fingerprint and applies the permutations with
spmmdb_permute_csr_f32 / spmmdb_permute_rows_f32.  Rabbit order: unavailable in this image (its demo needs
boost, libnuma and tcmalloc, none installed, no network).
"""
import os
import subprocess
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from spmm_denseblock_b200 import graphs  # noqa: E402

M1, M2, M3 = -7046029254386353131, -4417276706812531889, 1609587929392839161  # odd 64-bit multipliers


def _mix(x, m):
    x = x * m
    return x ^ (x >> 31)


def signatures(n, src, dst, rounds=4):
    """Colour refinement on a directed graph: a vertex's colour after r rounds depends only on the isomorphism
    class of its r-hop in/out neighbourhood."""
    src, dst = torch.as_tensor(src, dtype=torch.int64), torch.as_tensor(dst, dtype=torch.int64)
    outdeg = torch.bincount(src, minlength=n)
    indeg = torch.bincount(dst, minlength=n)
    sig = _mix(outdeg * 1000003 + indeg, M1)
    for _ in range(rounds):
        a = torch.zeros(n, dtype=torch.int64).index_add_(0, src, _mix(sig[dst], M2))
        b = torch.zeros(n, dtype=torch.int64).index_add_(0, dst, _mix(sig[src], M3))
        sig = _mix(sig, M1) + _mix(a, M2) + _mix(b, M3)
    return sig


def edges_of(rp, ci):
    n = rp.size - 1
    src = np.repeat(np.arange(n, dtype=np.int64), np.diff(rp))
    return src, ci.astype(np.int64)


def recover_permutation(n, old_edges, new_edges):
    so, sn = signatures(n, *old_edges), signatures(n, *new_edges)
    assert torch.unique(so).numel() == n, "signatures not unique: more refinement rounds needed"
    oo, on = torch.argsort(so), torch.argsort(sn)
    assert torch.equal(so[oo], sn[on]), "graphs are not isomorphic under the signatures"
    old2new = torch.empty(n, dtype=torch.int64)
    old2new[oo] = on
    # exact verification: the permuted edge set equals the emitted one
    ko = torch.sort(old2new[torch.as_tensor(old_edges[0])] * n + old2new[torch.as_tensor(old_edges[1])]).values
    kn = torch.sort(torch.as_tensor(new_edges[0]) * n + torch.as_tensor(new_edges[1])).values
    assert torch.equal(ko, kn), "recovered permutation does not reproduce the relabelled graph"
    return old2new.numpy().astype(np.int32)


def block_metrics(n, src, dst, bs):
    nb = (n + bs - 1) // bs
    nnzb = np.unique((src // bs) * nb + dst // bs).size
    return nnzb, src.size / (nnzb * bs * bs)


def fingerprint(rp, ci):
    return np.array([rp.size - 1, ci.size, int(rp[::997].astype(np.int64).sum() % (1 << 31)),
                     int((ci[::9973].astype(np.int64) * 31 % 1000003).sum() % (1 << 31))], np.int64)


def main():
    from oracle import oracle as O
    O.build()
    scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
    t = time.time()
    rp_t, ci_t = graphs.make_config_graph("c3_proteins", scale=scale, shuffle=True)
    rp, ci = rp_t.numpy(), ci_t.numpy()
    n = rp.size - 1
    src, dst = edges_of(rp, ci)
    print(f"graph: n={n} nnz={ci.size} ({time.time() - t:.1f} s)", flush=True)
    out = {"fingerprint": fingerprint(rp, ci), "scale": np.array([scale])}
    rows = []

    def report(name, s, d):
        for bs in (16, 32):
            nnzb, fill = block_metrics(n, s, d, bs)
            rows.append((name, bs, nnzb, fill))
            print(f"  {name:10s} bs={bs:2d} nnzb={nnzb:9d} fill={fill:.4f}", flush=True)

    report("shuffled", src, dst)
    # ideal = the planted order (the shuffle undone): upper bound for any reordering
    rp_i, ci_i = graphs.make_config_graph("c3_proteins", scale=scale, shuffle=False)
    report("ideal", *edges_of(rp_i.numpy(), ci_i.numpy()))

    t = time.time()
    nrp, nci = O.ref_rcmk_graph(rp, ci)
    print(f"reverseCuthillMcKee (reference, CPU): {time.time() - t:.1f} s", flush=True)
    rs, rd = edges_of(nrp, nci)
    out["old2new_rcm"] = recover_permutation(n, (src, dst), (rs, rd))
    report("rcm", rs, rd)

    gorder = "/root/reference/Gorder/Gorder"
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "c3.txt")
        t = time.time()
        np.savetxt(path, np.stack([src, dst], 1), fmt="%d")
        r = subprocess.run([gorder, path], cwd=tmp, capture_output=True, text=True)
        print(r.stdout.strip().replace("\n", " | "), f"({time.time() - t:.1f} s incl. text I/O)", flush=True)
        g = np.loadtxt(os.path.join(tmp, "c3_Gorder.txt"), dtype=np.int64, skiprows=1)
    gs, gd = g[:, 0], g[:, 1]
    # Gorder drops self loops (Graph.cpp readGraph); compare against the input without them
    keep = src != dst
    p = recover_permutation(n, (src[keep], dst[keep]), (gs, gd))
    out["old2new_gorder"] = p
    report("gorder", p[src], p[dst])

    out["metrics"] = np.array([(r[1], r[2], r[3]) for r in rows], np.float64)
    out["metric_names"] = np.array([r[0] for r in rows])
    dst_path = os.path.join(ROOT, "tests", "golden", "c3_orders.npz" if scale == 1.0 else f"c3_orders_{scale}.npz")
    np.savez_compressed(dst_path, **out)
    print("wrote", dst_path, os.path.getsize(dst_path), "bytes")


if __name__ == "__main__":
    main()

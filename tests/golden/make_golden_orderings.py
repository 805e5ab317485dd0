"""Generate tests/golden/ref_orderings.npz by RUNNING THE REFERENCE'S OWN ORDERINGS (BFSTraversal, reverseCuthillMcKee,
maxDegreeSort, reorder_strategy.cc:41-98, through oracle/_ref/libref_host.so built from /root/reference) on seeded
graphs.  The reference returns the REORDERED GRAPH, so that is what the fixture holds.  Run in the build container
(needs /root/reference); the .npz is committed and is what travels to the GPU box.

  python tests/golden/make_golden_orderings.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import random_csr, undirected_graph  # noqa: E402
from oracle import oracle as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    O.build()
    assert O.ref_host_lib() is not None and hasattr(O.ref_host_lib(), "ref_bfs_graph"), "oracle/_ref not built"
    rng = np.random.default_rng(4321)
    graphs = {
        # undirected: 11 components, isolated vertices, a hub, self loops, duplicated entries
        "und": undirected_graph(rng, 2500, 4, components=11, isolated=(0, 3, 2499), hubs={7: 600}, self_loops=12,
                                duplicates=9),
        # undirected with unsorted adjacency lists (BFSTraversal follows list order)
        "unsorted": undirected_graph(rng, 1200, 5, components=2, shuffle_rows=True),
        # directed: most vertices are reached from an earlier root's territory
        "directed": random_csr(rng, 2000, 2000, 1.5, hubs={1999: 700}, empty_rows={0, 1})[:2],
        # maximum degree <= 16: the reference's std::sort by degree is an insertion sort (stable), RCM is pinned exactly
        "lowdeg": undirected_graph(rng, 3000, 3, components=4, isolated=(5,)),
    }
    assert np.diff(graphs["lowdeg"][0]).max() <= 16
    out = {}
    for name, (rp, ci) in graphs.items():
        out[name + "_rowptr"], out[name + "_colind"] = rp, ci
        out[name + "_bfs_rowptr"], out[name + "_bfs_colind"] = O.ref_bfs_graph(rp, ci)
        out[name + "_rcm_rowptr"], out[name + "_rcm_colind"] = O.ref_rcmk_graph(rp, ci)
        out[name + "_deg_rowptr"], out[name + "_deg_colind"] = O.ref_max_degree_sort_graph(rp, ci)
    np.savez_compressed(os.path.join(HERE, "ref_orderings.npz"), **out)
    print("written", os.path.join(HERE, "ref_orderings.npz"))


if __name__ == "__main__":
    main()

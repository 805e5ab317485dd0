"""The oracle (oracle/spmm_oracle.c) against every known-answer vector the reference's
own programs hold for the hot path, against outputs of the reference's own CPU code
(tests/golden/*.npz, made by tests/golden/make_golden.py from oracle/_ref), and against
scipy as an independent cross-check.  CPU only."""
import os

import numpy as np
import pytest
import scipy.sparse as sp

from helpers import random_csr


def test_g1_csrmm_cu_doc_example(oracle):
    # csrmm.cu:47-99 -> printed y (column-major)
    A = sp.csr_matrix(np.array([[1, 0, 2, 3], [0, 4, 0, 0], [5, 0, 6, 7], [0, 8, 0, 9]], np.float32))
    B = np.array([10, 20, 30, 40, 50, 60, 70, 80], np.float32)  # column-major 4x2
    out = oracle.csrmm(A.indptr, A.indices, A.data, B, 2, ldb=4, b_layout=oracle.COL_MAJOR)
    assert out.T.ravel().tolist() == [190, 80, 510, 520, 430, 240, 1230, 1200]


def test_g2_bsrmm_cu(oracle):
    # bsrmm.cu:50-61: bs 2, row-major blocks, B column-major 6x2 = 1..12
    out = oracle.bsrmm([0, 2, 4], [0, 2, 1, 2], [0, 4, 2, 7, 1, 8, 2, 0, 9, 0, 0, 2, 0, 6, 7, 0], 2,
                       np.arange(1, 13, dtype=np.float32), 2, 6, trans_b=0)
    assert out.T.ravel().tolist() == [61, 26, 63, 43, 139, 92, 153, 97]


def test_g3_block_cublas(oracle):
    # block_cublas.cu:49-56,87-90: column-major blocks, row-major Y, column-major z, beta 1 on zeros
    out = oracle.block_gemm([0, 2, 3], [0, 1, 1], [0, 3, 1, 2, 4, 2, 0, 0, 0, 5, 0, 8], 2,
                            [6, 0, 0, 0, 7, 5, 4, 3, 0, 0, 0, 7], 3, np.zeros(12))
    assert out.T.ravel().tolist() == [16, 26, 0, 20, 19, 20, 0, 15, 5, 10, 0, 56]


def test_g4_spmm_cc_small(oracle):
    # spmm.cc:45-52
    out = oracle.spmm_cc_csr([0, 1, 3], [1, 0, 2], [[3, 9, 2], [4, 6, 7], [5, 8, 1]])
    assert out.tolist() == [[4, 6, 7], [8, 17, 3]]
    if oracle.ref_spmm_lib() is not None:  # the reference's own csr_spmm
        assert oracle.ref_csr_spmm([0, 1, 3], [1, 0, 2], [[3, 9, 2], [4, 6, 7], [5, 8, 1]]).tolist() == out.tolist()


def test_g6_divide_matrix(oracle):
    # divide.cu:138-153 (commented-out test_divide_matrix): n 4, bs 2, density 0.6
    crp, cci, brp, bci, bv = oracle.divide_matrix([0, 1, 2, 4, 7], [1, 2, 1, 3, 0, 1, 2], 2, 0.6)
    assert crp.tolist() == [0, 1, 2, 3, 4] and cci.tolist() == [1, 2, 3, 2]
    assert brp.tolist() == [0, 0, 1] and bci.tolist() == [0] and bv.ravel().tolist() == [0, 1, 1, 1]


def test_golden_ref_csr_spmm(oracle, golden_dir):
    d = np.load(os.path.join(golden_dir, "ref_csr_spmm.npz"))
    out = oracle.spmm_cc_csr(d["rowptr"], d["colind"], d["dense"])
    assert np.array_equal(out, d["out"])  # same loop order, same doubles -> identical
    # the fp32 value-carrying oracle agrees with it on unit weights
    o32 = oracle.csrmm(d["rowptr"], d["colind"], np.ones(d["colind"].size, np.float32),
                       d["dense"].astype(np.float32), d["dense"].shape[1])
    assert np.allclose(o32, d["out"], rtol=0, atol=1e-4)


def test_golden_ref_calculate_nnzb(oracle, golden_dir):
    d = np.load(os.path.join(golden_dir, "ref_calculate_nnzb.npz"))
    n = d["rowptr"].size - 1
    for bs, want in zip(d["bs"], d["nnzb"]):
        assert oracle.calculate_nnzb(d["rowptr"], d["colind"], int(bs)) == int(want)
        brp, bci, bv = oracle.csr2bsr(d["rowptr"], d["colind"], np.ones(d["colind"].size, np.float32), n, n, int(bs))
        assert bci.size == int(want) and brp[-1] == int(want)


def test_golden_ref_permutate(oracle, golden_dir):
    d = np.load(os.path.join(golden_dir, "ref_permutate.npz"))
    rp, ci = oracle.permutate_csr(d["rowptr"], d["colind"], d["old2new"])
    assert np.array_equal(rp, d["p_rowptr"]) and np.array_equal(ci, d["p_colind"])


@pytest.mark.parametrize("bs", [2, 3, 4, 16, 32])
@pytest.mark.parametrize("shape", [(64, 64), (100, 90), (37, 130)])
def test_csr2bsr_against_scipy(oracle, bs, shape):
    rng = np.random.default_rng(bs * 1000 + shape[0])
    m, n = shape
    rp, ci, v = random_csr(rng, m, n, 5, hubs={1: min(n, 60)}, empty_rows={0, 7})
    brp, bci, bv = oracle.csr2bsr(rp, ci, v, m, n, bs)
    mp, np_ = -(-m // bs) * bs, -(-n // bs) * bs
    A = sp.csr_matrix((v, ci, rp), shape=(m, n))
    A.resize((mp, np_))
    S = A.tobsr((bs, bs))
    S.sort_indices()  # scipy does not sort block indices
    # scipy drops nothing either: same pattern of touched blocks
    assert np.array_equal(brp, S.indptr) and np.array_equal(bci, S.indices)
    assert np.array_equal(bv, S.data.astype(np.float32))
    # column-major blocks are the transposed tiles
    _, _, bvc = oracle.csr2bsr(rp, ci, v, m, n, bs, direction=1)
    assert np.array_equal(bvc, bv.transpose(0, 2, 1))


@pytest.mark.parametrize("bs", [2, 5, 16])
def test_bsr2csr_closed_form_and_roundtrip(oracle, bs):
    rng = np.random.default_rng(7 + bs)
    m = n = 12 * bs
    rp, ci, v = random_csr(rng, m, n, 4, empty_rows={3})
    brp, bci, bv = oracle.csr2bsr(rp, ci, v, m, n, bs)
    crp, cci, cv = oracle.bsr2csr(brp, bci, bv, bs)
    nnzb = bci.size
    assert cci.size == nnzb * bs * bs and crp[-1] == nnzb * bs * bs  # bsr2csr.cu:177
    cnt = np.diff(brp)
    for i in range(m // bs):
        for r in range(bs):
            assert crp[i * bs + r] == brp[i] * bs * bs + r * cnt[i] * bs
    # same matrix (explicit zeros kept), ascending columns inside each row
    A = sp.csr_matrix((v, ci, rp), shape=(m, n)).toarray()
    B = sp.csr_matrix((cv, cci, crp), shape=(m, n)).toarray()
    assert np.array_equal(A, B)
    for i in range(m):
        assert np.all(np.diff(cci[crp[i]:crp[i + 1]]) > 0)
    # column-major blocks expand to the same CSR
    _, _, bvc = oracle.csr2bsr(rp, ci, v, m, n, bs, direction=1)
    crp2, cci2, cv2 = oracle.bsr2csr(brp, bci, bvc, bs, direction=1)
    assert np.array_equal(crp, crp2) and np.array_equal(cci, cci2) and np.array_equal(cv, cv2)


def test_check_result_style_agreement(oracle):
    # check_result.cu:99-261 at a size the oracle finishes: csr2bsr(bs 4) + bsrmm == csrmm, |diff| < 1e-4
    rng = np.random.default_rng(1234)
    m = n = 512
    rp, ci, v = random_csr(rng, m, n, 5)
    dim = 64
    y = np.where(np.arange(n * dim) % 2 == 0, 0.5, -0.5).astype(np.float32)  # constDenseMatrix :50-61
    z1 = oracle.csrmm(rp, ci, v, y, dim, ldb=dim, b_layout=oracle.ROW_MAJOR)
    brp, bci, bv = oracle.csr2bsr(rp, ci, v, m, n, 4)
    z2 = oracle.bsrmm(brp, bci, bv, 4, y, dim, dim, trans_b=1)
    assert np.max(np.abs(z1 - z2)) < 1e-4
    # and the fp32 sequential-FMA restatement of the reference kernel is within 1e-5 relative
    z3 = oracle.csrmm_f32_seq(rp, ci, v, y, dim)
    scale = oracle.csrmm(rp, ci, np.abs(v), np.abs(y), dim)
    assert np.max(np.abs(z3 - z1) / np.maximum(scale, 1e-30)) < 1e-5


def test_permutate_properties(oracle):
    rng = np.random.default_rng(5)
    n = 200
    rp, ci, _ = random_csr(rng, n, n, 6)
    perm = rng.permutation(n).astype(np.int32)
    prp, pci = oracle.permutate_csr(rp, ci, perm)
    A = sp.csr_matrix((np.ones(ci.size), ci, rp), shape=(n, n)).toarray()
    P = np.zeros((n, n))
    P[perm, np.arange(n)] = 1  # new = P old
    want = P @ A @ P.T
    got = sp.csr_matrix((np.ones(pci.size), pci, prp), shape=(n, n)).toarray()
    assert np.array_equal(want, got)
    inv = np.argsort(perm).astype(np.int32)
    brp, bci = oracle.permutate_csr(prp, pci, inv)
    assert np.array_equal(brp, rp) and np.array_equal(bci, ci)


def test_golden_ref_divide_matrix(oracle, golden_dir):
    """oracle.divide_matrix against outputs of the reference's OWN divide_matrix (divide.cu:52-127, compiled
    from /root/reference into oracle/_ref by oracle/Makefile; fixture made by tests/golden/make_golden.py)."""
    d = np.load(os.path.join(golden_dir, "ref_divide.npz"))
    for k, (bs, dens) in enumerate(d["cases"]):
        crp, cci, brp, bci, bv = oracle.divide_matrix(d["rowptr"], d["colind"], int(bs), float(dens))
        assert np.array_equal(crp, d[f"c{k}_csr_rowptr"]) and np.array_equal(cci, d[f"c{k}_csr_colind"])
        assert np.array_equal(brp, d[f"c{k}_bsr_rowptr"]) and np.array_equal(bci, d[f"c{k}_bsr_colind"])
        assert np.array_equal(bv, d[f"c{k}_bsr_val"])
        # every non-zero lands in exactly one part
        assert cci.size + int(np.count_nonzero(bv)) == d["colind"].size
        if oracle.ref_divide_lib() is not None:  # live, when oracle/_ref is present
            live = oracle.ref_divide_matrix(d["rowptr"], d["colind"], int(bs), float(dens))
            for a, b in zip(live, (crp, cci, brp, bci, bv)):
                assert np.array_equal(a, b)


def test_divide_threshold_matches_reference_expression():
    """spmmdb_divide_threshold (host function of the product library, no CUDA) = the smallest count the
    reference's `float occupy = (count * 1.0) / bnum; occupy >= density` accepts (divide.cu:89-93)."""
    import spmm_denseblock_b200 as pkg
    rng = np.random.default_rng(5)
    for bs in (2, 3, 8, 16, 32, 64):
        for dens in [0.6, 0.1, 0.5, 1.0, 0.999, 1e-4, 0.3333333] + rng.uniform(0, 1, 20).tolist():
            dens = float(np.float32(dens))
            thr = pkg.divide_threshold(bs, dens)
            ok = lambda c: np.float32((c * 1.0) / (bs * bs)) >= np.float32(dens)  # noqa: E731
            assert ok(thr) and thr >= 1 and (thr == 1 or not ok(thr - 1)), (bs, dens, thr)


def test_block_density_dist_restatement_matches_the_reference(oracle):
    """calculate_block_density_dist (block_density_dist.cpp:47-86): the numpy restatement against the reference's
    own function compiled from /root/reference (oracle/_ref/libref_host.so) on seeded inputs."""
    if oracle.ref_host_lib() is None or not hasattr(oracle.ref_host_lib(), "ref_block_density_dist"):
        pytest.skip("oracle/_ref/libref_host.so not built")
    from helpers import random_csr
    rng = np.random.default_rng(5)
    for n, avg in ((600, 30), (257, 90), (64, 50)):
        rp, ci, _ = random_csr(rng, n, n, avg)
        for bs in (2, 3, 4, 8, 16, 32):
            assert np.array_equal(oracle.block_density_dist(rp, ci, bs), oracle.ref_block_density_dist(rp, ci, bs))


def test_bfs_and_degree_order_restatements_match_the_reference(oracle):
    """oracle.bfs_order / degree_order against the reference's own BFSTraversal / maxDegreeSort
    (reorder_strategy.cc:41-98, compiled into oracle/_ref/libref_host.so).  The reference returns the reordered
    graph, so the restated permutation is applied with the (golden-pinned) permutate restatement and the graphs
    are compared.  maxDegreeSort leaves equal degrees in std::sort's order: row lengths always agree, the rows
    themselves when all degrees differ."""
    if oracle.ref_host_lib() is None or not hasattr(oracle.ref_host_lib(), "ref_bfs_graph"):
        pytest.skip("oracle/_ref/libref_host.so not built")
    from helpers import random_csr, undirected_graph
    rng = np.random.default_rng(11)
    cases = [undirected_graph(rng, 500, 4, components=7, isolated=(0, 17, 499), hubs={3: 200}, self_loops=5,
                              duplicates=6),
             undirected_graph(rng, 300, 3, components=1, shuffle_rows=True),
             random_csr(rng, 400, 400, 5)[:2],      # not symmetric: the sequential search does not care
             (np.zeros(6, np.int32), np.zeros(0, np.int32))]
    for rp, ci in cases:
        n = rp.size - 1
        p = oracle.bfs_order(rp, ci)
        assert np.array_equal(np.sort(p), np.arange(n))
        got, want = oracle.permutate_csr(rp, ci, p), oracle.ref_bfs_graph(rp, ci)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
        q = oracle.degree_order(rp)
        assert np.array_equal(np.sort(q), np.arange(n))
        got, want = oracle.permutate_csr(rp, ci, q), oracle.ref_max_degree_sort_graph(rp, ci)
        assert np.array_equal(got[0], want[0])
    # all degrees distinct: vertex i lists i of the other vertices
    n = 120
    perm = rng.permutation(n)
    rows = [np.sort(rng.choice(n, size=d, replace=False)) for d in perm]
    rp = np.zeros(n + 1, np.int32)
    rp[1:] = np.cumsum(perm)
    ci = np.concatenate(rows).astype(np.int32)
    got, want = oracle.permutate_csr(rp, ci, oracle.degree_order(rp)), oracle.ref_max_degree_sort_graph(rp, ci)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])


def test_rcm_order_restatement_matches_the_reference_where_std_sort_is_stable(oracle):
    """reverseCuthillMcKee (reorder_strategy.cc:57-66) sorts every list by degree with std::sort, which is an insertion
    sort (stable) for lists of at most 16 entries: on graphs of maximum degree <= 16 the restatement (stable sort +
    BFSTraversal) reproduces the reference's output graph bit for bit."""
    if oracle.ref_host_lib() is None:
        pytest.skip("oracle/_ref/libref_host.so not built")
    from helpers import undirected_graph
    rng = np.random.default_rng(12)
    for n, avg, comps in ((4000, 3, 5), (1500, 2.2, 40), (600, 4, 1)):
        rp, ci = undirected_graph(rng, n, avg, components=comps, isolated=(5,), shuffle_rows=True)
        assert np.diff(rp).max() <= 16
        p = oracle.rcm_order(rp, ci)
        got, want = oracle.permutate_csr(rp, ci, p), oracle.ref_rcmk_graph(rp, ci)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])


def test_golden_ref_orderings(oracle, golden_dir):
    """The restated orderings against the committed outputs of the reference's own BFSTraversal /
    reverseCuthillMcKee / maxDegreeSort (tests/golden/ref_orderings.npz, made by make_golden_orderings.py): BFS graphs
    bit for bit on every case (undirected, unsorted lists, directed); RCM bit for bit where the reference's sort is
    stable (maximum degree <= 16); the degree order's row lengths always."""
    d = np.load(os.path.join(golden_dir, "ref_orderings.npz"))
    for name in ("und", "unsorted", "directed", "lowdeg"):
        rp, ci = d[name + "_rowptr"], d[name + "_colind"]
        got = oracle.permutate_csr(rp, ci, oracle.bfs_order(rp, ci))
        assert np.array_equal(got[0], d[name + "_bfs_rowptr"]) and np.array_equal(got[1], d[name + "_bfs_colind"]), name
        got = oracle.permutate_csr(rp, ci, oracle.degree_order(rp))
        assert np.array_equal(got[0], d[name + "_deg_rowptr"]), name
        got = oracle.permutate_csr(rp, ci, oracle.rcm_order(rp, ci))
        assert np.array_equal(np.sort(np.diff(got[0])), np.sort(np.diff(d[name + "_rcm_rowptr"]))), name
        if name == "lowdeg":
            assert np.array_equal(got[0], d[name + "_rcm_rowptr"]) and np.array_equal(got[1], d[name + "_rcm_colind"])

"""Graph files: the library's host-side readers / writers against the reference's text
formats (golden bytes written by the reference's own dumpCSRToFile, its loaders when oracle/_ref is present,
and the numpy restatement), the binary cache format, and the error behaviour.  No GPU needed."""
import os
import subprocess

import numpy as np
import pytest
import torch

import spmm_denseblock_b200 as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a, np.int32))


def _random_graph(rng, n, deg):
    counts = rng.poisson(deg, n)
    counts[rng.integers(0, n, 3)] = 0  # empty rows
    rowptr = np.zeros(n + 1, np.int32)
    rowptr[1:] = np.cumsum(counts)
    colind = np.concatenate([np.sort(rng.choice(n, c, replace=False)) for c in counts] + [np.zeros(0, np.int64)])
    return rowptr, colind.astype(np.int32)


def test_golden_text_bytes_and_g7_edge_list(golden_dir, tmp_path):
    d = np.load(os.path.join(golden_dir, "ref_text_formats.npz"))
    p = str(tmp_path / "toy")
    S.write_graph(p, _t(d["rowptr"]), _t(d["colind"]), S.GRAPH_TEXT)
    assert open(p + "_indptr.txt").read() == str(d["indptr_txt"])    # the reference's bytes
    assert open(p + "_indices.txt").read() == str(d["indices_txt"])
    rp, ci = S.read_graph(p, S.GRAPH_TEXT)
    assert np.array_equal(rp.numpy(), d["rowptr"]) and np.array_equal(ci.numpy(), d["colind"])
    g = str(tmp_path / "g.txt")
    open(g, "w").write(str(d["edge_txt"]))
    rp, ci = S.read_graph(g, S.GRAPH_EDGELIST)
    assert np.array_equal(rp.numpy(), d["g_rowptr"]) and np.array_equal(ci.numpy(), d["g_colind"])
    assert rp.tolist() == [0, 2, 4, 6, 7, 8, 10] and ci.tolist() == [1, 5, 0, 2, 1, 3, 2, 5, 0, 4]


@pytest.mark.parametrize("threads", ["1", "7"])
def test_text_and_binary_round_trip_against_oracle(oracle, tmp_path, monkeypatch, threads):
    monkeypatch.setenv("SPMMDB_IO_THREADS", threads)
    rng = np.random.default_rng(5)
    rowptr, colind = _random_graph(rng, 60000, 9)  # ~3.5 MB of text: several parser ranges
    p, q = str(tmp_path / "a"), str(tmp_path / "b")
    S.write_graph(p, _t(rowptr), _t(colind), S.GRAPH_TEXT)
    oracle.dump_csr_text(q, rowptr, colind)
    for name in ("_indptr.txt", "_indices.txt"):
        assert open(p + name, "rb").read() == open(q + name, "rb").read()
    orp, oci = oracle.load_csr_text(p)
    rp, ci = S.read_graph(p)  # AUTO -> text, no binary file yet
    assert np.array_equal(rp.numpy(), orp) and np.array_equal(ci.numpy(), oci)
    assert np.array_equal(orp, rowptr) and np.array_equal(oci, colind)
    S.write_graph(p, rp, ci, S.GRAPH_BINARY)
    assert os.path.getsize(p + ".csrbin") == 32 + 4 * (rowptr.size + colind.size)
    os.remove(p + "_indices.txt")  # AUTO must now take the binary file
    rp2, ci2 = S.read_graph(p)
    assert torch.equal(rp2, rp) and torch.equal(ci2, ci)


def test_against_the_reference_loaders(oracle, tmp_path):
    if oracle.ref_host_lib() is None:
        pytest.skip("oracle/_ref/libref_host.so not built")
    rng = np.random.default_rng(6)
    rowptr, colind = _random_graph(rng, 4000, 6)
    p = str(tmp_path / "r")
    oracle.ref_dump_csr(p, rowptr, colind)         # written by the reference
    rp, ci = S.read_graph(p, S.GRAPH_TEXT)
    assert np.array_equal(rp.numpy(), rowptr) and np.array_equal(ci.numpy(), colind)
    q = str(tmp_path / "w")
    S.write_graph(q, rp, ci, S.GRAPH_TEXT)         # read back by the reference
    for name in ("_indptr.txt", "_indices.txt"):
        assert open(p + name, "rb").read() == open(q + name, "rb").read()
    rrp, rci = oracle.ref_load_csr(q, rowptr.size, colind.size)
    assert np.array_equal(rrp, rowptr) and np.array_equal(rci, colind)
    # edge list: shuffled pairs with duplicates, one pair per line / several per line / tabs
    n, e = 500, 6000
    x, y = rng.integers(0, n, e), rng.integers(0, n, e)
    x[:50], y[:50] = x[50:100], y[50:100]          # duplicate edges stay (load_data.cc:179-182)
    g = str(tmp_path / "g.txt")
    with open(g, "w") as f:
        f.write(f"{n} {e}\n")
        for k in range(e):
            f.write(f"{x[k]}\t{y[k]}" + ("\n" if k % 3 else "  "))
    rp, ci = S.read_graph(g, S.GRAPH_EDGELIST)
    rrp, rci = oracle.ref_load_graph_csr(g, n + 1, e)
    orp, oci = oracle.load_edge_list_csr(g)
    assert np.array_equal(rp.numpy(), rrp) and np.array_equal(ci.numpy(), rci)
    assert np.array_equal(orp, rrp) and np.array_equal(oci, rci)
    assert ci.numel() == e


def test_edge_list_header_nnz_wins_and_empty_graph(oracle, tmp_path):
    g = str(tmp_path / "g.txt")
    open(g, "w").write("4 3\n2 1\n0 3\n2 0\n3 3\n")  # a 4th pair beyond the header count is ignored
    rp, ci = S.read_graph(g, S.GRAPH_EDGELIST)
    orp, oci = oracle.load_edge_list_csr(g)
    assert rp.tolist() == orp.tolist() == [0, 1, 1, 3, 3] and ci.tolist() == oci.tolist() == [3, 0, 1]
    p = str(tmp_path / "e")
    S.write_graph(p, _t([0, 0, 0]), _t([]), S.GRAPH_TEXT)
    assert open(p + "_indices.txt").read() == "0\n\n"
    rp, ci = S.read_graph(p)
    assert rp.tolist() == [0, 0, 0] and ci.numel() == 0
    S.write_graph(p, rp, ci)
    rp, ci = S.read_graph(p, S.GRAPH_BINARY)
    assert rp.tolist() == [0, 0, 0] and ci.numel() == 0


def test_errors_are_loud(tmp_path):
    p = str(tmp_path / "x")
    with pytest.raises(S.SpmmdbError, match="IO"):
        S.read_graph(p)                                        # missing
    open(p + "_indptr.txt", "w").write("4\n0 1 2 ")            # short
    open(p + "_indices.txt", "w").write("3\n0 1 2 \n")
    with pytest.raises(S.SpmmdbError, match="IO"):
        S.read_graph(p)
    open(p + "_indptr.txt", "w").write("4\n0 2 1 3 \n")        # not monotone
    with pytest.raises(S.SpmmdbError, match="INVALID_VALUE"):
        S.read_graph(p)
    open(p + "_indptr.txt", "w").write("4\n0 1 2 4 \n")        # rowptr[n] != nnz
    with pytest.raises(S.SpmmdbError, match="INVALID_VALUE"):
        S.read_graph(p)
    open(p + ".csrbin", "wb").write(b"NOTAGRAPH" * 8)          # bad magic
    with pytest.raises(S.SpmmdbError, match="IO"):
        S.read_graph(p)
    S.write_graph(p, _t([0, 1, 2]), _t([1, 0]))
    with open(p + ".csrbin", "r+b") as f:                      # truncated payload
        f.truncate(40)
    with pytest.raises(S.SpmmdbError, match="IO"):
        S.read_graph(p, S.GRAPH_BINARY)
    g = str(tmp_path / "g.txt")
    open(g, "w").write("3 2\n0 1\n5 0\n")                       # source node out of range
    with pytest.raises(S.SpmmdbError, match="INVALID_VALUE"):
        S.read_graph(g, S.GRAPH_EDGELIST)
    # a column id outside [0, n) would become an out-of-bounds gather on the device: refused by every reader
    for bad in ("3 2\n0 1\n1 3\n", "3 2\n0 1\n1 -1\n"):
        open(g, "w").write(bad)
        with pytest.raises(S.SpmmdbError, match="INVALID_VALUE"):
            S.read_graph(g, S.GRAPH_EDGELIST)
    q = str(tmp_path / "y")
    open(q + "_indptr.txt", "w").write("4\n0 1 2 3 \n")
    open(q + "_indices.txt", "w").write("3\n0 3 1 \n")        # column 3 of a 3 x 3 matrix
    with pytest.raises(S.SpmmdbError, match="INVALID_VALUE"):
        S.read_graph(q, S.GRAPH_TEXT)
    open(q + "_indices.txt", "w").write("3\n0 2 1 \n")
    S.read_graph(q, S.GRAPH_TEXT)
    import numpy as np
    with open(q + ".csrbin", "wb") as f:
        f.write(b"SPMMDBG1" + np.array([3, 3, 3], np.int64).tobytes() + np.array([0, 1, 2, 3], np.int32).tobytes()
                + np.array([0, -2, 1], np.int32).tobytes())
    with pytest.raises(S.SpmmdbError, match="INVALID_VALUE"):
        S.read_graph(q, S.GRAPH_BINARY)


def test_graph2bin_driver(tmp_path):
    import __graft_entry__ as g
    exe = os.path.join(ROOT, "drivers", "bin", "graph2bin")
    if not os.path.exists(exe):
        g.build()
    rng = np.random.default_rng(8)
    rowptr, colind = _random_graph(rng, 3000, 5)
    p = str(tmp_path / "toy")
    S.write_graph(p, _t(rowptr), _t(colind), S.GRAPH_TEXT)
    r = subprocess.run([exe, p], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and f"n = 3000 nnz = {colind.size}" in r.stdout, r.stdout + r.stderr
    rp, ci = S.read_graph(p, S.GRAPH_BINARY)
    assert np.array_equal(rp.numpy(), rowptr) and np.array_equal(ci.numpy(), colind)
    e = str(tmp_path / "edges.txt")
    rows = np.repeat(np.arange(3000), np.diff(rowptr))
    order = rng.permutation(colind.size)
    with open(e, "w") as f:
        f.write(f"3000 {colind.size}\n" + "".join(f"{rows[k]} {colind[k]}\n" for k in order))
    q = str(tmp_path / "fromedges")
    r = subprocess.run([exe, "--edges", e, q], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    for name in ("_indptr.txt", "_indices.txt"):
        assert open(p + name, "rb").read() == open(q + name, "rb").read()
    assert subprocess.run([exe, str(tmp_path / "nope")], capture_output=True).returncode == 1

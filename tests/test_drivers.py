"""The drop-in drivers (drivers/*.cu) keep the reference's command lines, file formats and printed
lines and agree with the vendor library where the reference compared against it."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "drivers", "bin")
PROGS = ["run_csrmm", "run_bsrmm", "test_csrmm", "test_bsrmm", "check_result", "csr2bsr", "bsr2csr", "block_cublas", "divide", "graph2bin"]


def _build():
    import __graft_entry__ as g
    if not all(os.path.exists(os.path.join(BIN, p)) for p in PROGS):
        g.build()


def test_drivers_build_and_link():
    _build()
    for p in PROGS:
        exe = os.path.join(BIN, p)
        assert os.access(exe, os.X_OK), exe
        out = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
        assert "libspmmdb.so" in out and "not found" not in out.split("libspmmdb.so")[1].splitlines()[0]


def _run(args, cwd, env=None):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run(args, cwd=cwd, capture_output=True, text=True, timeout=600, env=e)
    return r.returncode, r.stdout + r.stderr


def _checksum(out):
    return float(re.search(r"checksum: (\S+)", out).group(1))


@pytest.fixture(scope="module")
def graph_dir(tmp_path_factory):
    """tmp/<graph>_indptr.txt / _indices.txt in the reference's text format (load_data.cc:125-141)."""
    from helpers import random_csr
    d = tmp_path_factory.mktemp("drv")
    os.makedirs(d / "tmp")
    rng = np.random.default_rng(9)
    rp, ci, _ = random_csr(rng, 3000, 3000, 11, hubs={3: 1200}, values="unit")
    with open(d / "tmp" / "toy_indptr.txt", "w") as f:
        f.write(f"{rp.size}\n" + " ".join(map(str, rp.tolist())))
    with open(d / "tmp" / "toy_indices.txt", "w") as f:
        f.write(f"{ci.size}\n" + " ".join(map(str, ci.tolist())))
    return str(d), rp.size - 1, ci.size


@pytest.mark.gpu
def test_run_csrmm_cli_and_agreement(graph_dir):
    _build()
    cwd, n, nnz = graph_dir
    sums = {}
    for impl, tb in (("gespmm", 1), ("cusparseScsrmm2", 1), ("cusparseScsrmm2", 0), ("cusparseScsrmm", 0),
                     ("cusparseSpMM", 1)):
        rc, out = _run([os.path.join(BIN, "run_csrmm"), "toy", "128", impl, str(tb)], cwd)
        assert rc == 0, out
        assert f"graph = toy dim = 128 csrmmImpl = {impl} transposeB = {tb}" in out
        assert f"n={n} nnz={nnz}" in out
        assert len(re.findall(r"csrmm cost time:  \d+\.\d{10} ms", out)) == 10
        assert re.search(r"average csrmm cost time: \d+\.\d{10} ms", out)
        sums[(impl, tb)] = _checksum(out)
    # same seeded B for every run: row-major runs agree with each other and with the vendor library
    base = sums[("gespmm", 1)]
    assert abs(sums[("cusparseScsrmm2", 1)] - base) <= 1e-3 * max(1.0, abs(base))
    assert abs(sums[("cusparseSpMM", 1)] - base) <= 1e-3 * max(1.0, abs(base))
    assert abs(sums[("cusparseScsrmm2", 0)] - sums[("cusparseScsrmm", 0)]) <= 1e-6 * max(1.0, abs(base))


@pytest.mark.gpu
def test_run_csrmm_reads_the_binary_cache(graph_dir, tmp_path):
    """drivers/graph2bin writes tmp/<graph>.csrbin; the drivers then load it instead of parsing the text pair
    (SPMMDB_GRAPH_AUTO) and print the same result."""
    import shutil
    _build()
    cwd, n, nnz = graph_dir
    work = tmp_path / "w"
    shutil.copytree(os.path.join(cwd, "tmp"), work / "tmp")
    rc, out_text = _run([os.path.join(BIN, "run_csrmm"), "toy", "128", "gespmm", "1"], str(work))
    assert rc == 0, out_text
    rc, out = _run([os.path.join(BIN, "graph2bin"), "tmp/toy"], str(work))
    assert rc == 0 and f"n = {n} nnz = {nnz}" in out, out
    os.remove(work / "tmp" / "toy_indices.txt")  # only the binary file can serve the graph now
    rc, out_bin = _run([os.path.join(BIN, "run_csrmm"), "toy", "128", "gespmm", "1"], str(work))
    assert rc == 0, out_bin
    assert f"n={n} nnz={nnz}" in out_bin
    assert _checksum(out_bin) == _checksum(out_text)
    os.remove(work / "tmp" / "toy.csrbin")
    rc, out = _run([os.path.join(BIN, "run_csrmm"), "toy", "128", "gespmm", "1"], str(work))
    assert rc != 0  # missing graph files fail loudly instead of running on garbage


@pytest.mark.gpu
def test_run_bsrmm_cli_and_agreement(graph_dir):
    _build()
    cwd, n, nnz = graph_dir
    for bs in (4, 16):
        sums = {}
        for impl, math in (("cusparse", "fp32"), ("rocsparse", "fp32"), ("rocsparse", "tf32")):
            rc, out = _run([os.path.join(BIN, "run_bsrmm"), "toy", str(bs), "64", impl], cwd, {"SPMMDB_BSRMM_MATH": math})
            assert rc == 0, out
            assert "tmp/toy" in out and f"n={n} nnz={nnz}" in out
            assert re.search(r"numVal = \d+", out) and re.search(r"density:  \d+\.\d{10}", out)
            assert re.search(r"bsrmm cost time:  \d+\.\d{10} ms", out)
            sums[(impl, math)] = _checksum(out)
        ref = sums[("cusparse", "fp32")]
        assert abs(sums[("rocsparse", "fp32")] - ref) <= 1e-3 * max(1.0, abs(ref))
        assert abs(sums[("rocsparse", "tf32")] - ref) <= 2e-2 * max(1.0, abs(ref)) + 1.0


@pytest.mark.gpu
def test_check_result_and_conversion_drivers(tmp_path):
    _build()
    rc, out = _run([os.path.join(BIN, "check_result"), "8192", "0.01", "64"], str(tmp_path))
    assert rc == 0 and "all checks passed" in out, out
    rc, out = _run([os.path.join(BIN, "csr2bsr"), "0.1"], str(tmp_path))
    assert rc == 0 and "PASSED" in out, out
    rc, out = _run([os.path.join(BIN, "bsr2csr"), "0.05", "4"], str(tmp_path))
    assert rc == 0 and "PASSED" in out, out
    for math in ("fp32", "tf32"):
        rc, out = _run([os.path.join(BIN, "block_cublas")], str(tmp_path), {"SPMMDB_BSRMM_MATH": math})
        assert rc == 0 and "PASSED" in out, out


@pytest.mark.gpu
def test_microbench_drivers_print_reference_lines(tmp_path):
    _build()
    rc, out = _run([os.path.join(BIN, "test_csrmm"), "0.0001", "64", "gespmm"], str(tmp_path))
    assert rc == 0 and re.search(r"csrmm cost time: \d+\.\d{10} ms\nGFLOPs: \d+\.\d{10}", out), out
    s1 = _checksum(out)
    rc, out = _run([os.path.join(BIN, "test_csrmm"), "0.0001", "64", "cusparse"], str(tmp_path))
    assert rc == 0, out
    assert abs(_checksum(out) - s1) <= 1e-3 * max(1.0, abs(s1))
    for tb in (0, 1):
        sums = []
        for impl in ("rocsparse", "cusparse"):
            rc, out = _run([os.path.join(BIN, "test_bsrmm"), "0.0005", "16", "64", impl, str(tb)], str(tmp_path),
                           {"SPMMDB_BSRMM_MATH": "fp32"})
            assert rc == 0 and re.search(r"bsrmm cost time: \d+\.\d{10} ms\nGFLOPs: \d+\.\d{10}", out), out
            sums.append(_checksum(out))
        assert abs(sums[0] - sums[1]) <= 1e-3 * max(1.0, abs(sums[1]))
        # the drop-in default arithmetic is the tcgen05 tf32 kernel (stated 1e-2)
        rc, out = _run([os.path.join(BIN, "test_bsrmm"), "0.0005", "16", "64", "rocsparse", str(tb)], str(tmp_path))
        assert rc == 0, out
        assert abs(_checksum(out) - sums[1]) <= 2e-2 * max(1.0, abs(sums[1])) + 1.0
    rc, out = _run([os.path.join(BIN, "test_block_cublas"), "0.0005", "16", "64"], str(tmp_path))
    assert rc == 0 and "block gemm cost time" in out, out


@pytest.mark.gpu
def test_divide_cli_and_agreement(tmp_path):
    """divide <graph> <bsize> <dim> <rocsparse|cusparse> <transB> <density> (divide.cu:205-391): printed lines
    kept, hybrid product == plain CSR product, our BSR part == the vendor cusparseSbsrmm on the same split."""
    _build()
    os.makedirs(tmp_path / "tmp")
    rng = np.random.default_rng(21)
    n, band = 2000, 16
    rows = []
    for i in range(n):
        c = set(rng.choice(n, size=int(rng.poisson(5)), replace=False).tolist())
        if (i // band) % 2 == 0:
            lo = (i // band) * band
            c |= set(j for j in range(lo, min(lo + band, n)) if rng.uniform() < 0.7)
        rows.append(sorted(c))
    rp = np.concatenate([[0], np.cumsum([len(r) for r in rows])])
    ci = np.concatenate([np.asarray(r, np.int64) for r in rows])
    with open(tmp_path / "tmp" / "band_indptr.txt", "w") as f:
        f.write(f"{rp.size}\n" + " ".join(map(str, rp.tolist())))
    with open(tmp_path / "tmp" / "band_indices.txt", "w") as f:
        f.write(f"{ci.size}\n" + " ".join(map(str, ci.tolist())))
    sums = {}
    for impl, math, tb in (("cusparse", "fp32", 1), ("rocsparse", "fp32", 1), ("rocsparse", "tf32", 1),
                           ("rocsparse", "fp32", 0), ("cusparse", "fp32", 0)):
        rc, out = _run([os.path.join(BIN, "divide"), "band", "16", "64", impl, str(tb), "0.4"], str(tmp_path),
                       {"SPMMDB_BSRMM_MATH": math})
        assert rc == 0, out
        m = re.search(r"csr nnz = (\d+)    bsr nnzb = (\d+)", out)
        assert m and int(m.group(1)) < ci.size and int(m.group(2)) > 0, out
        for line in ("tmp/band", "csr to adj...", "divide matrix...", "prepare y and z...", "end"):
            assert line in out, out
        assert re.search(r"csrmm cost time:  \d+\.\d{10} ms", out) and re.search(r"bsrmm cost time:  \d+\.\d{10} ms", out)
        assert re.search(r"total cost time:  \d+\.\d{10} ms", out) and re.search(r"\d+\.\d{5}\+\d+\.\d{5}=\d+\.\d{5}", out)
        diff = float(re.search(r"max \|hybrid - csr\| = (\S+),", out).group(1))
        assert diff <= (2e-2 if math == "tf32" else 1e-4), out
        sums[(impl, math, tb)] = float(re.search(r"checksum (\S+)", out).group(1))
    for tb in (0, 1):
        ref = sums[("cusparse", "fp32", tb)]
        assert abs(sums[("rocsparse", "fp32", tb)] - ref) <= 1e-3 * max(1.0, abs(ref))


@pytest.mark.gpu
def test_known_answer_programs(tmp_path):
    """csrmm / bsrmm without arguments (csrmm.cu:26-218, bsrmm.cu:35-177): the reference's example matrices, its
    printed format, goldens G1 / G2 of SURVEY 8c."""
    _build()
    rc, out = _run([os.path.join(BIN, "csrmm")], str(tmp_path))
    assert rc == 0 and "testing examples" in out and out.rstrip().endswith("end"), out
    assert "190.000000 80.000000 510.000000 520.000000 430.000000 240.000000 1230.000000 1200.000000" in out
    rc, out = _run([os.path.join(BIN, "bsrmm")], str(tmp_path))
    assert rc == 0, out
    assert out.count("61.000000 26.000000 63.000000 43.000000 139.000000 92.000000 153.000000 97.000000") == 2, out

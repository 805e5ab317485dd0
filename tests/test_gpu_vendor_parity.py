"""Parity against what the reference itself runs on this path, on the same device arrays:
  * cuSPARSE csr2bsrNnz + csr2bsr (run_bsrmm.cu:121-142) and bsr2csr (bsr2csr.cu:186-188): all arrays BIT-EQUAL;
  * cusparseSbsrmm (run_bsrmm.cu:161) and cusparseSpMM (successor of the csrmm calls, run_csrmm.cu:133-142);
  * the cublasSgemm-per-block loop (block_cublas.cu:123-136);
  * the reference's OWN kernels compiled for sm_100a from /root/reference where they lie
    (baseline/_ref/libref_gpu.so: gespmm_csrmm.h spmm_test2, rocsparse_bsrmm.h).
The comparators live in baseline/ (test infrastructure, never linked by the product)."""
import os
import sys

import numpy as np
import pytest
import torch

from helpers import random_bsr, random_csr

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline"))


@pytest.fixture(scope="module")
def pkg():
    import spmm_denseblock_b200 as p
    assert torch.cuda.is_available()
    p.lib()
    return p


@pytest.fixture(scope="module")
def cmp():
    import compare
    compare.vendor()
    return compare


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(params=["block_row", "warp_merge"])
def conversion_path(request, monkeypatch):
    monkeypatch.setenv("SPMMDB_CONVERT_BR", "1" if request.param == "block_row" else "0")
    return request.param


@pytest.mark.parametrize("bs", [2, 3, 4, 8, 16, 32, 64])
@pytest.mark.parametrize("direction", [0, 1])
@pytest.mark.parametrize("shape", [(1000, 1200, 9), (257, 4099, 40), (33, 31, 3)])  # ragged last block row / column
def test_csr2bsr_bit_equal_to_cusparse(pkg, cmp, conversion_path, bs, direction, shape):
    m, n, avg = shape
    rng = np.random.default_rng(bs * 13 + direction + m)
    rp, ci, v = random_csr(rng, m, n, avg, hubs={1: min(n, 900)}, empty_rows={0, m - 1})
    drp, dci, dv = dev(rp), dev(ci), dev(v)
    brp, bci, bv = pkg.csr2bsr(m, n, drp, dci, dv, bs, direction)
    try:
        vrp, vci, vv = cmp.cusparse_csr2bsr(m, n, drp, dci, dv, bs, direction)
    except RuntimeError as e:  # e.g. CUSPARSE_STATUS_INVALID_VALUE for some (shape, blockDim) pairs in 12.5
        pytest.skip(f"cuSPARSE rejects this case: {e}")
    assert torch.equal(brp, vrp) and torch.equal(bci, vci)
    assert torch.equal(bv.view(torch.int32), vv.view(torch.int32))  # bit pattern, zeros included


@pytest.mark.parametrize("bs", [2, 3, 8, 16, 32, 64])
@pytest.mark.parametrize("direction", [0, 1])
def test_bsr2csr_bit_equal_to_cusparse(pkg, cmp, bs, direction):
    rng = np.random.default_rng(bs * 17 + direction)
    mb, nb = 37, 53
    brp, bci, bv = random_bsr(rng, mb, nb, bs, 6, values="sparse")
    d = [dev(x) for x in (brp, bci, bv)]
    ours = pkg.bsr2csr(mb, nb, *d, bs, direction)
    theirs = cmp.cusparse_bsr2csr(mb, nb, *d, bs, direction)
    for o, t in zip(ours, theirs):
        assert torch.equal(o.view(torch.int32), t.view(torch.int32))


def test_conversion_round_trip_through_both_libraries(pkg, cmp, conversion_path):
    """csr2bsr (ours) -> bsr2csr (cuSPARSE) -> csr2bsr (cuSPARSE) -> bsr2csr (ours): the two BSR matrices and the
    two expanded CSR matrices are the same bits."""
    rng = np.random.default_rng(4)
    m = n = 640
    rp, ci, v = random_csr(rng, m, n, 12)
    for bs in (16, 32):
        b1 = pkg.csr2bsr(m, n, dev(rp), dev(ci), dev(v), bs)
        c1 = cmp.cusparse_bsr2csr(m // bs, n // bs, *b1, bs)
        b2 = cmp.cusparse_csr2bsr(m, n, *c1, bs)
        c2 = pkg.bsr2csr(m // bs, n // bs, *b2, bs)
        for x, y in zip(b1, b2):
            assert torch.equal(x, y)
        for x, y in zip(c1, c2):
            assert torch.equal(x, y)


@pytest.mark.parametrize("bs", [16, 32])
@pytest.mark.parametrize("math,tol", [(0, 1e-5), (1, 1e-2), (2, 1e-2)])
def test_block_gemm_tensor_core_shapes_vs_oracle_and_cublas_loop(pkg, cmp, oracle, bs, math, tol):
    """spmmdb_block_gemm_f32 (column-major blocks, row-major Y, column-major Z, alpha = beta = 1 accumulate) at
    the block sizes the tcgen05 kernel takes, against the oracle's block_cublas.cu:123-136 restatement and
    against the cublasSgemm loop itself."""
    rng = np.random.default_rng(bs + math)
    mb, dim = 40, 128
    brp, bci, bv = random_bsr(rng, mb, mb, bs, 5)
    m = mb * bs
    Y = rng.uniform(-1, 1, (m, dim)).astype(np.float32)
    Z0 = rng.uniform(-1, 1, (dim, m)).astype(np.float32)  # column-major z with ld m, pre-filled: beta = 1
    Z = dev(Z0)
    pkg.block_gemm(mb, dim, bci.size, bs, dev(brp), dev(bci), dev(bv), dev(Y), Z, math=math)
    got = Z.cpu().numpy().T
    ref = oracle.block_gemm(brp, bci, bv, bs, Y, dim, Z0)
    absA = np.abs(bv)
    scale = oracle.block_gemm(brp, bci, absA, bs, np.abs(Y), dim, np.abs(Z0))
    assert float(np.max(np.abs(got - ref) / np.maximum(scale, 1e-30))) <= tol
    Zc = dev(Z0)
    cmp.cublas_block_gemm_loop(mb, dim, bs, brp.astype(np.int32), bci.astype(np.int32), dev(bv), dev(Y), Zc)
    gotc = Zc.cpu().numpy().T
    assert float(np.max(np.abs(gotc - ref) / np.maximum(scale, 1e-30))) <= 1e-3  # cuBLAS may use tf32-free fp32
    assert float(np.max(np.abs(got - gotc) / np.maximum(scale, 1e-30))) <= max(tol, 1e-3)


@pytest.mark.parametrize("bs,trans_b,direction", [(16, 1, 0), (32, 1, 0), (16, 0, 0), (32, 0, 1), (5, 1, 0)])
def test_bsrmm_fp32_vs_cusparse_bsrmm(pkg, cmp, bs, trans_b, direction):
    rng = np.random.default_rng(bs * 3 + trans_b)
    mb, kb, n = 50, 60, 96
    brp, bci, bv = random_bsr(rng, mb, kb, bs, 7)
    m, k = mb * bs, kb * bs
    B = rng.uniform(-1, 1, (k, n) if trans_b else (n, k)).astype(np.float32)
    ldb = n if trans_b else k
    C0 = rng.uniform(-1, 1, (n, m)).astype(np.float32)
    d = [dev(x) for x in (bv, brp, bci)]
    ours, theirs = dev(C0), dev(C0)
    pkg.bsrmm(direction, trans_b, mb, n, kb, bci.size, 0.5, *d, bs, dev(B), ldb, 2.0, ours, m, math=0)
    cmp.cusparse_bsrmm(direction, trans_b, mb, n, kb, bci.size, 0.5, *d, bs, dev(B), ldb, 2.0, theirs, m)
    assert float((ours - theirs).abs().max()) <= 1e-4  # the reference's own bar (check_result.cu:233-246)


@pytest.mark.parametrize("K", [64, 128, 256, 300])
def test_csrmm_bits_of_the_reference_gespmm_kernel(pkg, cmp, K):
    """The reference's spmm_test2 (gespmm_csrmm.h:95-166, compiled for sm_100a) and spmmdb_csrmm_f32 run the
    same sequential fp32 FMA chain per output element: rows of at most one segment are BIT-EQUAL on the box;
    longer rows agree to 1e-5 relative."""
    if cmp.ref_gpu() is None:
        pytest.skip("baseline/_ref/libref_gpu.so not built (needs /root/reference at build time)")
    rng = np.random.default_rng(K)
    m, k = 5000, 7000
    rp, ci, v = random_csr(rng, m, k, 11, hubs={3: 3000, 4000: 700}, empty_rows={0, 9})
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    d = [dev(x) for x in (rp, ci, v, B)]
    ours = torch.full((m, K), float("nan"), device="cuda")
    theirs = torch.full((m, K), float("nan"), device="cuda")
    pkg.gespmm_csrmm(m, K, *d, ours)
    cmp.ref_gespmm(m, K, *d, theirs)
    torch.cuda.synchronize()
    short = torch.from_numpy(np.diff(rp) <= pkg.csrmm_segment(ci.size, m=m)).cuda()
    assert torch.equal(ours[short], theirs[short])
    absref = torch.zeros(m, K, device="cuda")
    cmp.ref_gespmm(m, K, d[0], d[1], d[2].abs(), d[3].abs(), absref)
    assert float(((ours - theirs).abs() / absref.clamp_min(1e-30)).max()) <= 1e-5


def test_reference_rocsparse_kernel_vs_ours(pkg, cmp, oracle):
    """the rocSPARSE-port kernels derive the output column from blockDim.y (rocsparse_bsrmm_impl.h:
    27, 326, 413), so for n <= 256 they leave C unwritten; bsrmmnt_small (bs = 2, op(B) = T) is unaffected and
    must agree with us.  Recorded here on the box; our result is the mathematical product (oracle)."""
    if cmp.ref_gpu() is None:
        pytest.skip("baseline/_ref/libref_gpu.so not built")
    rng = np.random.default_rng(9)
    # (1) the unaffected kernel: bs = 2, trans_b = T
    mb, kb, n, bs = 300, 280, 64, 2
    brp, bci, bv = random_bsr(rng, mb, kb, bs, 6)
    B = rng.uniform(-1, 1, (kb * bs, n)).astype(np.float32)
    d = [dev(x) for x in (bv, brp, bci)]
    ours = torch.zeros(n, mb * bs, device="cuda")
    theirs = torch.zeros(n, mb * bs, device="cuda")
    pkg.bsrmm(0, 1, mb, n, kb, bci.size, 1.0, *d, bs, dev(B), n, 0.0, ours, mb * bs, math=0)
    cmp.ref_rocsparse_bsrmm(0, 1, mb, n, kb, bci.size, 1.0, *d, bs, dev(B), n, 0.0, theirs, mb * bs)
    torch.cuda.synchronize()
    assert float((ours - theirs).abs().max()) <= 1e-4
    # (2) the affected kernel: bs = 16, n = 64 -> the reference writes nothing; ours equals the oracle
    mb, kb, n, bs = 40, 40, 64, 16
    brp, bci, bv = random_bsr(rng, mb, kb, bs, 5)
    B = rng.uniform(-1, 1, (kb * bs, n)).astype(np.float32)
    d = [dev(x) for x in (bv, brp, bci)]
    theirs = torch.full((n, mb * bs), float("nan"), device="cuda")
    cmp.ref_rocsparse_bsrmm(0, 1, mb, n, kb, bci.size, 1.0, *d, bs, dev(B), n, 0.0, theirs, mb * bs)
    torch.cuda.synchronize()
    written = float((~torch.isnan(theirs)).float().mean())
    print(f"reference rocsparse_bsrmm_template bs=16 n=64: fraction of C written = {written}")
    assert written == 0.0  # the output-column bug, as SURVEY 8 a4 derived from the source
    ours = torch.zeros(n, mb * bs, device="cuda")
    pkg.bsrmm(0, 1, mb, n, kb, bci.size, 1.0, *d, bs, dev(B), n, 0.0, ours, mb * bs, math=0)
    ref = oracle.bsrmm(brp, bci, bv, bs, B, n, n, 1)
    assert np.max(np.abs(ours.cpu().numpy().T - ref)) <= 1e-4


@pytest.mark.parametrize("K,layouts", [(128, (0, 0)), (48, (1, 1)), (64, (0, 1))])
def test_csrmm_vs_cusparse_spmm(pkg, cmp, K, layouts):
    """cusparseSpMM is what the reference's cusparseScsrmm / csrmm2 call sites (run_csrmm.cu:133-142) become on a
    current toolkit; same layouts, alpha / beta."""
    b_layout, c_layout = layouts
    rng = np.random.default_rng(K)
    m, k = 3000, 2500
    rp, ci, v = random_csr(rng, m, k, 9, hubs={7: 2000})
    B = rng.uniform(-1, 1, (K, k) if b_layout else (k, K)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (K, m) if c_layout else (m, K)).astype(np.float32)
    d = [dev(x) for x in (rp, ci, v, B)]
    ours, theirs = dev(C0), dev(C0)
    pkg.csrmm(*d, ours, m, k, K, b_layout=b_layout, c_layout=c_layout, alpha=1.5, beta=-0.5)
    cmp.cusparse_spmm(m, k, K, *d, theirs, alg=0, b_layout=b_layout, c_layout=c_layout, alpha=1.5, beta=-0.5)
    torch.cuda.synchronize()
    assert float((ours - theirs).abs().max()) <= 2e-4


@pytest.mark.parametrize("bs", [16, 32, 5])
def test_transpose_blocks_once_equals_direction_column(pkg, bs):
    """spmmdb_bsr_transpose_blocks_f32: column-major blocks converted once + SPMMDB_DIR_ROW give the bits of the
    SPMMDB_DIR_COLUMN call (which transposes per call on the tf32 path); twice = identity."""
    rng = np.random.default_rng(bs)
    mb, n = 30, 128
    brp, bci, bv = random_bsr(rng, mb, mb, bs, 5)
    d_rp, d_ci, d_v = dev(brp), dev(bci), dev(bv)
    rowmajor = pkg.bsr_transpose_blocks(d_v, bs)
    assert torch.equal(pkg.bsr_transpose_blocks(rowmajor, bs), d_v)
    want = torch.from_numpy(bv.reshape(-1, bs, bs).transpose(0, 2, 1).copy().reshape(-1)).cuda()
    assert torch.equal(rowmajor, want)
    B = dev(rng.uniform(-1, 1, (mb * bs, n)).astype(np.float32))
    for math in (0, 1):
        c1 = torch.zeros(n, mb * bs, device="cuda")
        c2 = torch.zeros(n, mb * bs, device="cuda")
        pkg.bsrmm(1, 1, mb, n, mb, bci.size, 1.0, d_v, d_rp, d_ci, bs, B, n, 0.0, c1, mb * bs, math=math)
        pkg.bsrmm(0, 1, mb, n, mb, bci.size, 1.0, rowmajor, d_rp, d_ci, bs, B, n, 0.0, c2, mb * bs, math=math)
        assert torch.equal(c1, c2), math

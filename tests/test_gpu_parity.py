"""GPU parity tests: every call goes through the C ABI (libspmmdb.so) and is compared with
the CPU oracle on the same seeded inputs.  Bars (BASELINE.json north_star):
  * csr2bsr / bsr2csr: bit-exact.
  * fp32 CSR path: relative error <= 1e-5 against the fp64 oracle (error measured against
    sum_j |a_ij b_jk|), and BIT-EXACT against the reference kernel's own fp32 FMA order for
    rows no longer than one segment.
  * fp32 BSR path: <= 1e-5; tf32/bf16 tensor-core BSR path: <= 1e-2.
"""
import numpy as np
import pytest
import torch

from helpers import random_bsr, random_csr

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-5
TOL_TC = 1e-2


@pytest.fixture(scope="module")
def pkg():
    import spmm_denseblock_b200 as p
    assert torch.cuda.is_available()
    p.lib()
    return p


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(autouse=True, params=["block_row", "warp_merge"])
def conversion_path(request, monkeypatch):
    """csr2bsr and the hybrid split have two device paths with identical outputs (one CTA per block row with
    shared-memory bitmap ranks; one warp per block row merging the rows, which also serves shapes whose bitmap
    does not fit shared memory).  Every test that converts runs on both."""
    name = request.node.name
    converts = any(k in name for k in ("csr2bsr", "divide", "conversion", "round_trip", "roundtrip", "check_result",
                                       "metrics"))
    if request.param == "warp_merge":
        if not converts:
            pytest.skip("no conversion in this test")
        monkeypatch.setenv("SPMMDB_CONVERT_BR", "0")
    elif converts:
        monkeypatch.setenv("SPMMDB_CONVERT_BR", "1")  # else small inputs would all pick the warp-merge path
    else:
        monkeypatch.delenv("SPMMDB_CONVERT_BR", raising=False)


def err_vs(oracle, got, ref64, rp, ci, v, Bm, K, **kw):
    scale = oracle.csrmm(rp, ci, np.abs(v), np.abs(Bm), K, **kw)
    return float(np.max(np.abs(got.astype(np.float64) - ref64) / np.maximum(scale, 1e-30)))


# ----------------------------------------------------------------------------- CSR
@pytest.mark.parametrize("K", [1, 7, 16, 32, 64, 100, 128, 256, 300, 516, 1030])
def test_csrmm_rowmajor_widths(pkg, oracle, K):
    rng = np.random.default_rng(K)
    m, k = 777, 555
    rp, ci, v = random_csr(rng, m, k, 9, hubs={5: 300, 700: 555}, empty_rows={0, 6, 776})
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    C = torch.full((m, K), float("nan"), device="cuda")
    pkg.gespmm_csrmm(m, K, dev(rp), dev(ci), dev(v), dev(B), C)
    got = C.cpu().numpy()
    ref = oracle.csrmm(rp, ci, v, B, K)
    assert err_vs(oracle, got, ref, rp, ci, v, B, K) <= TOL_FP32
    # rows within one segment follow the reference kernel's FMA order exactly
    seq = oracle.csrmm_f32_seq(rp, ci, v, B, K)
    short = np.diff(rp) <= pkg.csrmm_segment(ci.size, m=m)
    assert np.array_equal(got[short], seq[short])
    assert np.all(got[[0, 6, 776]] == 0)  # empty rows are written (gespmm_csrmm.h:133-135)


def test_csrmm_alpha_beta_and_ld(pkg, oracle):
    rng = np.random.default_rng(3)
    m, k, K, ldb, ldc = 300, 200, 64, 80, 72
    rp, ci, v = random_csr(rng, m, k, 6, hubs={9: 150})
    B = rng.uniform(-1, 1, (k, ldb)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (m, ldc)).astype(np.float32)
    C = dev(C0)
    pkg.csrmm(dev(rp), dev(ci), dev(v), dev(B), C, m, k, K, ldb=ldb, ldc=ldc, alpha=0.5, beta=-2.0)
    ref = oracle.csrmm(rp, ci, v, B, K, ldb=ldb, C0=C0, ldc=ldc, alpha=0.5, beta=-2.0)
    got = C.cpu().numpy()
    assert np.allclose(got[:, :K], ref, rtol=0, atol=2e-5)
    assert np.array_equal(got[:, K:], C0[:, K:])  # padding columns untouched


@pytest.mark.parametrize("b_layout,c_layout", [(1, 1), (0, 1), (1, 0)])
def test_csrmm_cusparse_layouts(pkg, oracle, b_layout, c_layout):
    # cusparseScsrmm (B, C column-major, run_csrmm.cu:135) and csrmm2+transB (B row-major, C column-major, :139)
    rng = np.random.default_rng(11 + b_layout * 2 + c_layout)
    m, k, K = 257, 131, 48
    rp, ci, v = random_csr(rng, m, k, 5)
    ldb = k + 3 if b_layout else K + 4
    ldc = m + 5 if c_layout else K + 4
    B = rng.uniform(-1, 1, (K if b_layout else k, ldb)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (K if c_layout else m, ldc)).astype(np.float32)
    C = dev(C0)
    pkg.csrmm(dev(rp), dev(ci), dev(v), dev(B), C, m, k, K, ldb=ldb, b_layout=b_layout, ldc=ldc, c_layout=c_layout,
              alpha=1.0, beta=1.0)
    ref = oracle.csrmm(rp, ci, v, B, K, ldb=ldb, b_layout=b_layout, C0=C0, ldc=ldc, c_layout=c_layout, alpha=1.0,
                       beta=1.0)
    got = C.cpu().numpy()
    got_mk = got[:, :m].T if c_layout else got[:, :K]
    assert np.allclose(got_mk, ref, rtol=0, atol=2e-5)


def test_csrmm_golden_g1_and_g4(pkg):
    # csrmm.cu:47-99 (column-major B and C)
    rp = np.array([0, 3, 4, 7, 9], np.int32)
    ci = np.array([0, 2, 3, 1, 0, 2, 3, 1, 3], np.int32)
    v = np.array([1, 2, 3, 4, 5, 6, 7, 8, 9], np.float32)
    B = np.array([10, 20, 30, 40, 50, 60, 70, 80], np.float32)
    C = torch.zeros(8, device="cuda")
    pkg.csrmm(dev(rp), dev(ci), dev(v), dev(B), C, 4, 4, 2, ldb=4, b_layout=1, ldc=4, c_layout=1)
    assert C.cpu().tolist() == [190, 80, 510, 520, 430, 240, 1230, 1200]
    # spmm.cc:45-52 (unit weights, row-major)
    C = torch.zeros(2, 3, device="cuda")
    pkg.gespmm_csrmm(2, 3, dev(np.array([0, 1, 3], np.int32)), dev(np.array([1, 0, 2], np.int32)),
                     dev(np.ones(3, np.float32)), dev(np.array([[3, 9, 2], [4, 6, 7], [5, 8, 1]], np.float32)), C)
    assert C.cpu().tolist() == [[4, 6, 7], [8, 17, 3]]


def test_csrmm_long_rows_are_sharding_independent(pkg, oracle):
    """Rows longer than a segment: deterministic, and identical whether the row is computed
    inside the full matrix or inside a row shard (SURVEY 8e verification)."""
    rng = np.random.default_rng(99)
    m, k, K = 400, 3000, 128
    rp, ci, v = random_csr(rng, m, k, 4, hubs={3: 2900, 150: 1000, 399: 65})
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    dB = dev(B)
    full = torch.empty(m, K, device="cuda")
    nnz = ci.size
    pkg.csrmm(dev(rp), dev(ci), dev(v), dB, full, m, k, K, nnz=nnz)
    again = torch.empty(m, K, device="cuda")
    pkg.csrmm(dev(rp), dev(ci), dev(v), dB, again, m, k, K, nnz=nnz)
    assert torch.equal(full, again)
    ref = oracle.csrmm(rp, ci, v, B, K)
    assert err_vs(oracle, full.cpu().numpy(), ref, rp, ci, v, B, K) <= TOL_FP32
    # shard rows [100, 400): rebased rowptr, same segment length as the full matrix
    lo = 100
    rps = (rp[lo:] - rp[lo]).astype(np.int32)
    cis, vs = ci[rp[lo]:], v[rp[lo]:]
    part = torch.empty(m - lo, K, device="cuda")
    pkg.set_csrmm_segment(pkg.csrmm_segment(nnz, m=m))  # the whole matrix's segment length
    try:
        pkg.csrmm(dev(rps), dev(cis), dev(vs), dB, part, m - lo, k, K)
    finally:
        pkg.set_csrmm_segment(0)
    assert torch.equal(part, full[lo:])


def test_csrmm_empty_and_degenerate(pkg):
    z = torch.zeros(1, dtype=torch.int32, device="cuda")
    C = torch.zeros(0, 8, device="cuda")
    pkg.csrmm(z, z[:0], torch.zeros(0, device="cuda"), torch.zeros(4, 8, device="cuda"), C, 0, 4, 8)
    # all-empty rows
    rp = torch.zeros(6, dtype=torch.int32, device="cuda")
    C = torch.full((5, 8), 3.0, device="cuda")
    pkg.csrmm(rp, z[:0], torch.zeros(0, device="cuda"), torch.ones(4, 8, device="cuda"), C, 5, 4, 8)
    assert torch.all(C == 0)


# ----------------------------------------------------------------------------- conversions
@pytest.mark.parametrize("bs", [2, 3, 4, 8, 16, 32, 48, 64, 130])
@pytest.mark.parametrize("direction", [0, 1])
def test_csr2bsr_bit_exact(pkg, oracle, bs, direction):
    rng = np.random.default_rng(bs * 7 + direction)
    m, n = 1000, 1200  # csr2bsr.cu:98-99 shape; ragged for most bs
    rp, ci, v = random_csr(rng, m, n, 8, hubs={4: 900, 999: 1200}, empty_rows={0, 1, 2, 3} if bs == 4 else {10})
    brp, bci, bv = pkg.csr2bsr(m, n, dev(rp), dev(ci), dev(v), bs, direction=direction)
    orp, oci, ov = oracle.csr2bsr(rp, ci, v, m, n, bs, direction=direction)
    assert np.array_equal(brp.cpu().numpy(), orp)
    assert np.array_equal(bci.cpu().numpy(), oci)
    assert np.array_equal(bv.cpu().numpy().view(np.uint32), ov.ravel().view(np.uint32))


@pytest.mark.parametrize("bs", [2, 5, 8, 16, 32, 64])
@pytest.mark.parametrize("direction", [0, 1])
def test_bsr2csr_bit_exact(pkg, oracle, bs, direction):
    rng = np.random.default_rng(bs + 100 * direction)
    mb, nb = 150, 170
    brp, bci, bv = random_bsr(rng, mb, nb, bs, 6, values="sparse")
    crp, cci, cv = pkg.bsr2csr(mb, nb, dev(brp), dev(bci), dev(bv), bs, direction=direction)
    orp, oci, ov = oracle.bsr2csr(brp, bci, bv, bs, direction=direction)
    assert np.array_equal(crp.cpu().numpy(), orp)
    assert np.array_equal(cci.cpu().numpy(), oci)
    assert np.array_equal(cv.cpu().numpy().view(np.uint32), ov.view(np.uint32))


@pytest.mark.parametrize("bs", [8, 16, 32])
def test_bsr2csr_block_runs_cross_empty_block_rows_and_unaligned_input(pkg, oracle, bs):
    """The float4 path walks bsr_rowptr along a warp's run of blocks: mostly-empty block rows, runs that span
    many block rows, and a bsr_val that is not 16-byte aligned (scalar path) all give the closed form."""
    rng = np.random.default_rng(40 + bs)
    mb, nb = 2500, 300
    brp, bci, bv = random_bsr(rng, mb, nb, bs, 0.6, values="sparse")
    assert (np.diff(brp) == 0).sum() > mb // 3
    orp, oci, ov = oracle.bsr2csr(brp, bci, bv, bs)
    crp, cci, cv = pkg.bsr2csr(mb, nb, dev(brp), dev(bci), dev(bv), bs)
    assert np.array_equal(crp.cpu().numpy(), orp) and np.array_equal(cci.cpu().numpy(), oci)
    assert np.array_equal(cv.cpu().numpy().view(np.uint32), ov.view(np.uint32))
    shifted = torch.empty(bv.size + 1, device="cuda")[1:]
    shifted.copy_(torch.from_numpy(bv.ravel()))
    assert shifted.data_ptr() % 16 != 0
    crp, cci, cv = pkg.bsr2csr(mb, nb, dev(brp), dev(bci), shifted, bs)
    assert np.array_equal(cci.cpu().numpy(), oci)
    assert np.array_equal(cv.cpu().numpy().view(np.uint32), ov.view(np.uint32))


def test_conversion_roundtrip_and_empty(pkg):
    rng = np.random.default_rng(5)
    m = n = 640
    rp, ci, v = random_csr(rng, m, n, 7)
    for bs in (4, 16, 32):
        brp, bci, bv = pkg.csr2bsr(m, n, dev(rp), dev(ci), dev(v), bs)
        crp, cci, cv = pkg.bsr2csr(m // bs, n // bs, brp, bci, bv, bs)
        keep = cv != 0  # v has no exact zeros: dropping the fill recovers the input
        assert torch.equal(cci[keep].cpu(), torch.from_numpy(ci))
        assert torch.equal(cv[keep].cpu(), torch.from_numpy(v))
    # empty matrix
    rp0 = torch.zeros(m + 1, dtype=torch.int32, device="cuda")
    brp, bci, bv = pkg.csr2bsr(m, n, rp0, rp0[:0], torch.zeros(0, device="cuda"), 16)
    assert bci.numel() == 0 and torch.all(brp == 0)


# ----------------------------------------------------------------------------- BSR
def bsr_case(rng, mb, nb, bs, n, trans_b, c_layout, avg=5):
    brp, bci, bv = random_bsr(rng, mb, nb, bs, avg)
    k, m = nb * bs, mb * bs
    ldb = (n + 4) if trans_b else (k + 8)
    ldc = (m + 8) if c_layout == 1 else (n + 4)
    B = rng.uniform(-1, 1, (k, ldb) if trans_b else (n, ldb)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (n, ldc) if c_layout == 1 else (m, ldc)).astype(np.float32)
    return brp, bci, bv, B, C0, ldb, ldc


def bsr_scale(oracle, brp, bci, bv, bs, B, n, ldb, trans_b, direction):
    return oracle.bsrmm(brp, bci, np.abs(bv), bs, np.abs(B), n, ldb, trans_b, direction=direction)


@pytest.mark.parametrize("bs", [2, 3, 4, 8, 16, 32, 40, 64])
@pytest.mark.parametrize("trans_b", [0, 1])
@pytest.mark.parametrize("direction", [0, 1])
def test_bsrmm_fp32(pkg, oracle, bs, trans_b, direction):
    rng = np.random.default_rng(bs * 4 + trans_b * 2 + direction)
    mb, nb, n = 37, 41, 70
    brp, bci, bv, B, C0, ldb, ldc = bsr_case(rng, mb, nb, bs, n, trans_b, 1)
    C = dev(C0)
    rc = pkg.rocsparse_bsrmm(direction, 0, trans_b, mb, n, nb, bci.size, 0.75, dev(bv), dev(brp), dev(bci), bs,
                             dev(B), ldb, -1.5, C, ldc)
    assert rc == 0
    ref = oracle.bsrmm(brp, bci, bv, bs, B, n, ldb, trans_b, direction=direction, alpha=0.75, beta=-1.5, C0=C0,
                       ldc=ldc)
    scale = 0.75 * bsr_scale(oracle, brp, bci, bv, bs, B, n, ldb, trans_b, direction) + 1.5 * np.abs(
        C0[:, :mb * bs].T)
    got = C.cpu().numpy()[:, :mb * bs].T
    assert np.max(np.abs(got - ref) / np.maximum(scale, 1e-30)) <= TOL_FP32
    assert np.array_equal(C.cpu().numpy()[:, mb * bs:], C0[:, mb * bs:])


def test_bsrmm_golden_g2_g3(pkg):
    # bsrmm.cu:50-61
    C = torch.zeros(8, device="cuda")
    pkg.bsrmm(0, 0, 2, 2, 3, 4, 1.0, dev(np.array([0, 4, 2, 7, 1, 8, 2, 0, 9, 0, 0, 2, 0, 6, 7, 0], np.float32)),
              dev(np.array([0, 2, 4], np.int32)), dev(np.array([0, 2, 1, 2], np.int32)), 2,
              dev(np.arange(1, 13, dtype=np.float32)), 6, 0.0, C, 4)
    assert C.cpu().tolist() == [61, 26, 63, 43, 139, 92, 153, 97]
    # block_cublas.cu:49-56,87-90
    Z = torch.zeros(12, device="cuda")
    pkg.block_gemm(2, 3, 3, 2, dev(np.array([0, 2, 3], np.int32)), dev(np.array([0, 1, 1], np.int32)),
                   dev(np.array([0, 3, 1, 2, 4, 2, 0, 0, 0, 5, 0, 8], np.float32)),
                   dev(np.array([6, 0, 0, 0, 7, 5, 4, 3, 0, 0, 0, 7], np.float32)), Z)
    assert Z.cpu().tolist() == [16, 26, 0, 20, 19, 20, 0, 15, 5, 10, 0, 56]


def test_bsrmm_status_codes(pkg):
    z = torch.zeros(4, device="cuda")
    zi = torch.zeros(4, dtype=torch.int32, device="cuda")
    assert pkg.rocsparse_bsrmm(0, 1, 0, 1, 1, 1, 1, 1.0, z, zi, zi, 2, z, 2, 0.0, z, 2) == -2  # trans_A != N
    assert pkg.rocsparse_bsrmm(0, 0, 0, 1, 1, 1, 1, 1.0, z, zi, zi, 2, z, 1, 0.0, z, 2) == -1  # ldb < kb*bs
    assert pkg.rocsparse_bsrmm(0, 0, 0, 1, 1, 1, 0, 1.0, z, zi, zi, 2, z, 2, 0.0, z, 2) == 0   # nnzb == 0


@pytest.mark.parametrize("math", [1, 2])
@pytest.mark.parametrize("bs", [16, 32, 64])
@pytest.mark.parametrize("direction", [0, 1])
@pytest.mark.parametrize("trans_b,c_layout,n", [(1, 0, 128), (1, 1, 128), (0, 1, 128), (1, 0, 300), (0, 1, 44)])
def test_bsrmm_tensor_core(pkg, oracle, math, bs, direction, trans_b, c_layout, n):
    """tcgen05 path (tf32 / bf16 inputs, fp32 accumulate): relative error <= 1e-2 (north_star),
    alpha/beta applied, every layout the reference interface has, ragged feature counts."""
    rng = np.random.default_rng(bs + math * 3 + trans_b + 7 * direction + n)
    mb, nb = 67, 59
    brp, bci, bv = random_bsr(rng, mb, nb, bs, 9)
    brp, bci, bv = _drop_row(brp, bci, bv, 4, bs)
    k, m = nb * bs, mb * bs
    ldb = n if trans_b else k
    ldc = m if c_layout == 1 else n
    B = rng.uniform(-1, 1, (k, ldb) if trans_b else (n, ldb)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (n, ldc) if c_layout == 1 else (m, ldc)).astype(np.float32)
    C = dev(C0)
    pkg.bsrmm(direction, trans_b, mb, n, nb, bci.size, 0.5, dev(bv), dev(brp), dev(bci), bs, dev(B), ldb, 2.0, C, ldc,
              c_layout=c_layout, math=math)
    ref = oracle.bsrmm(brp, bci, bv, bs, B, n, ldb, trans_b, direction=direction, alpha=0.5, beta=2.0, C0=C0, ldc=ldc,
                       c_layout=c_layout)
    scale = 0.5 * bsr_scale(oracle, brp, bci, bv, bs, B, n, ldb, trans_b, direction) + 2.0 * np.abs(
        C0.T if c_layout == 1 else C0)
    got = C.cpu().numpy()
    got = got.T if c_layout == 1 else got
    err = np.max(np.abs(got - ref) / np.maximum(scale, 1e-30))
    assert err <= TOL_TC, err
    # the tensor-core path must really have run: bf16 rounding is visible, fp32 would be ~1e-7
    if math == 2:
        assert err > 1e-5


def _drop_row(brp, bci, bv, row, bs):
    """Remove every block of block row `row` (an empty block row in the middle of a super-row)."""
    lo, hi = int(brp[row]), int(brp[row + 1])
    cnt = hi - lo
    bci = np.concatenate([bci[:lo], bci[hi:]])
    bv = np.concatenate([bv[:lo * bs * bs], bv[hi * bs * bs:]])
    brp = brp.copy()
    brp[row + 1:] -= cnt
    return brp, bci.astype(np.int32), bv.astype(np.float32)


def test_bsrmm_tensor_core_dense_strip_and_empty(pkg, oracle):
    """Super-rows whose block rows share every column block (one wide MMA per step), a fully empty
    super-row, and a ragged last super-row (mb not a multiple of the group size)."""
    rng = np.random.default_rng(77)
    bs, mb, nb, n = 16, 21, 12, 128
    rows = []
    for i in range(mb):
        if 8 <= i < 16:
            rows.append(np.zeros(0, np.int64))  # super-row 1 (block rows 8..15) is empty
        else:
            rows.append(np.arange(nb))           # dense strips
    brp = np.zeros(mb + 1, np.int32)
    brp[1:] = np.cumsum([len(r) for r in rows])
    bci = np.concatenate(rows).astype(np.int32)
    bv = rng.uniform(-1, 1, bci.size * bs * bs).astype(np.float32)
    B = rng.uniform(-1, 1, (nb * bs, n)).astype(np.float32)
    for math in (1, 2):
        C = torch.full((mb * bs, n), float("nan"), device="cuda")
        pkg.bsrmm(0, 1, mb, n, nb, bci.size, 1.0, dev(bv), dev(brp), dev(bci), bs, dev(B), n, 0.0, C, n, c_layout=0,
                  math=math)
        ref = oracle.bsrmm(brp, bci, bv, bs, B, n, n, 1)
        scale = bsr_scale(oracle, brp, bci, bv, bs, B, n, n, 1, 0)
        got = C.cpu().numpy()
        assert np.all(got[8 * bs:16 * bs] == 0)
        assert np.max(np.abs(got - ref) / np.maximum(scale, 1e-30)) <= TOL_TC


def test_check_result_harness(pkg, oracle):
    """check_result.cu:99-261 through the new engine: csr2bsr(bs 4) + bsrmm(N,T) == csrmm2(N,T),
    |z1 - z2| < 1e-4, on a Bernoulli(p) matrix with the +-0.5 checkerboard B."""
    from spmm_denseblock_b200 import graphs
    m = n = 4096
    rp_t, ci_t = graphs.bernoulli_csr(m, n, 0.01, seed=1234)
    rp, ci = rp_t.numpy(), ci_t.numpy()
    rng = np.random.default_rng(1234)
    v = rng.uniform(-1, 1, ci.size).astype(np.float32)
    dim = 64
    y = np.where(np.arange(n * dim) % 2 == 0, 0.5, -0.5).astype(np.float32)
    z1 = torch.zeros(dim, m, device="cuda")  # column-major m x dim
    pkg.csrmm(dev(rp), dev(ci), dev(v), dev(y), z1, m, n, dim, ldb=dim, b_layout=0, ldc=m, c_layout=1)
    brp, bci, bv = pkg.csr2bsr(m, n, dev(rp), dev(ci), dev(v), 4)
    z2 = torch.zeros(dim, m, device="cuda")
    pkg.bsrmm(0, 1, m // 4, dim, n // 4, bci.numel(), 1.0, bv, brp, bci, 4, dev(y), dim, 0.0, z2, m)
    assert float((z1 - z2).abs().max()) < 1e-4
    ref = oracle.csrmm(rp, ci, v, y, dim)
    assert np.max(np.abs(z1.cpu().numpy().T - ref)) < 1e-4


@pytest.mark.parametrize("K", [128, 100])
def test_csrmm_scratch_reuse_across_sizes(pkg, oracle, K):
    """The long-row queue lives in library scratch whose layout moves with nnz: a call must not trip over
    what an earlier call with another nnz left there (regression: `done` counters of a small matrix landed on
    the row records of a bigger one and its hub rows came out wrong)."""
    rng = np.random.default_rng(77)
    mats = []
    for m, avg, hubs in ((3000, 10, {3: 1500, 900: 700, 2999: 300}), (900, 9, {0: 400, 450: 200}),
                         (2000, 3, {5: 1999, 6: 1000, 7: 500, 1500: 90})):
        rp, ci, v = random_csr(rng, m, m, avg, hubs=hubs)
        B = rng.uniform(-1, 1, (m, K)).astype(np.float32)
        mats.append((m, rp, ci, v, B, oracle.csrmm(rp, ci, v, B, K)))
    for order in ((0, 1, 2), (2, 1, 0), (1, 0, 1), (0, 2, 0)):
        for i in order:
            m, rp, ci, v, B, ref = mats[i]
            C = torch.full((m, K), float("nan"), device="cuda")
            pkg.csrmm(dev(rp), dev(ci), dev(v), dev(B), C, m, m, K)
            assert err_vs(oracle, C.cpu().numpy(), ref, rp, ci, v, B, K) <= TOL_FP32, (order, i)


@pytest.mark.parametrize("K", [64, 128, 256, 300])
@pytest.mark.parametrize("alpha,beta", [(1.0, 0.0), (0.5, 2.0)])
def test_csrmm_understated_nnz_overflows_the_long_row_queue(pkg, oracle, K, alpha, beta):
    """The long-row queue is sized from the caller's nnz.  A caller that understates it (here: 64 for a matrix
    with four hub rows) overflows the queue; rows that do not fit are computed in place by the kernel that
    found them -- on the overlapped small-matrix path that is the queue-building pre-pass."""
    rng = np.random.default_rng(K)
    m = 2500
    rp, ci, v = random_csr(rng, m, m, 6, hubs={1: 2000, 700: 900, 701: 300, 2499: 130})
    B = rng.uniform(-1, 1, (m, K)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (m, K)).astype(np.float32)
    ref = oracle.csrmm(rp, ci, v, B, K, C0=C0, alpha=alpha, beta=beta)
    scale = abs(alpha) * oracle.csrmm(rp, ci, np.abs(v), np.abs(B), K) + abs(beta) * np.abs(C0)
    for nnz in (64, 200, None):
        C = dev(C0.copy())
        pkg.csrmm(dev(rp), dev(ci), dev(v), dev(B), C, m, m, K, alpha=alpha, beta=beta, nnz=nnz)
        err = np.max(np.abs(C.cpu().numpy().astype(np.float64) - ref) / np.maximum(scale, 1e-30))
        assert err <= TOL_FP32, (nnz, err)


@pytest.mark.parametrize("K,alpha,beta", [(256, 1.0, 0.0), (256, 0.5, 2.0), (128, 1.0, 0.0), (100, 1.0, 0.0), (192, 1.0, 0.0),
                                          (1024, 1.0, -1.0), (80, 1.0, 0.0)])
def test_csrmm_host_entry_slab_pipeline(pkg, oracle, K, alpha, beta):
    """spmmdb_csrmm_host_f32 (host buffers; feature slabs copied / computed / copied back in a pipeline when
    K allows) returns the bits of the device call on the same inputs."""
    import ctypes as C
    rng = np.random.default_rng(K)
    m = 5000
    rp, ci, v = random_csr(rng, m, m, 8, hubs={17: 3000, 4000: 700})
    B = rng.uniform(-1, 1, (m, K)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (m, K)).astype(np.float32)
    hC = C0.copy()
    L = pkg.lib()
    p = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
    rc = L.spmmdb_csrmm_host_f32(m, m, K, ci.size, p(rp), p(ci), p(v), p(B), p(hC), alpha, beta, C.c_void_p(0))
    assert rc == 0
    dC = dev(C0)
    pkg.csrmm(dev(rp), dev(ci), dev(v), dev(B), dC, m, m, K, alpha=alpha, beta=beta)
    assert np.array_equal(hC, dC.cpu().numpy())
    ref = oracle.csrmm(rp, ci, v, B, K, C0=C0, alpha=alpha, beta=beta)
    assert np.max(np.abs(hC - ref)) <= 1e-3


@pytest.mark.parametrize("bs", [16, 32, 64])
@pytest.mark.parametrize("direction", [0, 1])
def test_bsrmm_bf16_packed_blocks(pkg, oracle, bs, direction):
    """SPMMDB_MATH_BF16_PACKED: blocks converted once by spmmdb_bsr_pack_bf16; same result as the per-call bf16
    path bit for bit, within the tensor-core tolerance of the fp64 oracle; unsupported shapes are refused."""
    rng = np.random.default_rng(bs + direction)
    mb, nb, n = 37, 41, 128
    rp, ci, v = random_bsr(rng, mb, nb, bs, 6, values="sparse")
    B = rng.uniform(-1, 1, (nb * bs, n)).astype(np.float32)
    dv = dev(v)
    packed = pkg.bsr_pack_bf16(dv, bs)
    assert packed.dtype == torch.bfloat16 and torch.equal(packed.float(), dv.to(torch.bfloat16).float())
    outs = []
    for math, vals in ((pkg.MATH_BF16, dv), (pkg.MATH_BF16_PACKED, packed)):
        C = torch.full((mb * bs, n), float("nan"), device="cuda")
        pkg.bsrmm(direction, 1, mb, n, nb, ci.size, 1.0, vals, dev(rp), dev(ci), bs, dev(B), n, 0.0, C, n,
                  c_layout=pkg.ROW_MAJOR, math=math)
        outs.append(C)
    assert torch.equal(outs[0], outs[1])
    ref = oracle.bsrmm(rp, ci, v, bs, B, n, n, 1, direction=direction, c_layout=0)
    scale = oracle.bsrmm(rp, ci, np.abs(v), bs, np.abs(B), n, n, 1, direction=direction, c_layout=0)
    assert np.max(np.abs(outs[1].cpu().numpy() - ref) / np.maximum(scale, 1e-30)) <= TOL_TC
    # a block size the tensor-core kernel does not take cannot read bf16 blocks
    L = pkg.lib()
    z = torch.zeros(64, device="cuda")
    rc = L.spmmdb_bsrmm_f32(0, 1, 1, 8, 1, 1, 5, 1.0, z.data_ptr(), dev(np.array([0, 1], np.int32)).data_ptr(),
                            dev(np.array([0], np.int32)).data_ptr(), z.data_ptr(), 8, 0.0, z.data_ptr(), 8, 0, 3, None)
    assert rc == -2


# ----------------------------------------------------------------------------- divide (hybrid split)
def _banded_graph(rng, n, band=16, p_dense=0.7, avg=4):
    rows = []
    for i in range(n):
        c = set(rng.choice(n, size=min(int(rng.poisson(avg)), n), replace=False).tolist())
        if (i // band) % 2 == 0:
            lo = (i // band) * band
            c |= set(j for j in range(lo, min(lo + band, n)) if rng.uniform() < p_dense)
        rows.append(np.array(sorted(c), np.int32))
    rp = np.zeros(n + 1, np.int32)
    rp[1:] = np.cumsum([len(r) for r in rows])
    ci = np.concatenate(rows).astype(np.int32) if rp[-1] else np.zeros(0, np.int32)
    return rp, ci


def test_divide_golden_reference(pkg, golden_dir):
    """GPU divide against the outputs of the reference's own divide_matrix (unit weights, val = NULL)."""
    import os
    d = np.load(os.path.join(golden_dir, "ref_divide.npz"))
    rp, ci = d["rowptr"], d["colind"]
    n = rp.size - 1
    for k, (bs, dens) in enumerate(d["cases"]):
        bs = int(bs)
        crp, cci, cv, brp, bci, bv = pkg.divide(n, n, dev(rp), dev(ci), None, bs, float(dens))
        assert np.array_equal(crp.cpu().numpy(), d[f"c{k}_csr_rowptr"])
        assert np.array_equal(cci.cpu().numpy(), d[f"c{k}_csr_colind"])
        assert np.array_equal(brp.cpu().numpy(), d[f"c{k}_bsr_rowptr"])
        assert np.array_equal(bci.cpu().numpy(), d[f"c{k}_bsr_colind"])
        assert np.array_equal(bv.cpu().numpy(), d[f"c{k}_bsr_val"].ravel())
        assert np.all(cv.cpu().numpy() == 1.0)


@pytest.mark.parametrize("bs,dens", [(2, 0.6), (5, 0.3), (16, 0.25), (16, 0.6), (32, 0.1), (33, 0.05), (64, 0.02),
                                     (16, 0.0), (16, 1.5)])
@pytest.mark.parametrize("direction", [0, 1])
def test_divide_bit_exact_with_values(pkg, oracle, bs, dens, direction):
    rng = np.random.default_rng(bs * 7 + direction)
    n = 1000 + bs // 2  # ragged last block row
    rp, ci = _banded_graph(rng, n, band=max(bs, 4))
    v = rng.uniform(-1, 1, ci.size).astype(np.float32)
    crp, cci, cv, brp, bci, bv = pkg.divide(n, n, dev(rp), dev(ci), dev(v), bs, dens, direction=direction)
    # density <= 0: the reference would emit every (also empty) block; the product emits the non-empty ones
    o_dens = dens if dens > 0 else 1e-9
    ocrp, occi, obrp, obci, obv, ocv = oracle.divide_matrix(rp, ci, bs, o_dens, val=v)
    if direction == 1:
        obv = obv.transpose(0, 2, 1)
    assert np.array_equal(crp.cpu().numpy(), ocrp) and np.array_equal(cci.cpu().numpy(), occi)
    assert np.array_equal(cv.cpu().numpy(), ocv)
    assert np.array_equal(brp.cpu().numpy(), obrp) and np.array_equal(bci.cpu().numpy(), obci)
    assert np.array_equal(bv.cpu().numpy(), np.ascontiguousarray(obv).ravel())
    if dens <= 0:
        assert cci.numel() == 0      # everything is a block: plain csr2bsr
    if dens > 1:
        assert bci.numel() == 0      # nothing qualifies: the remainder is the input
        assert np.array_equal(cci.cpu().numpy(), ci)


def test_divide_empty_and_edge_inputs(pkg):
    z = torch.zeros(1, dtype=torch.int32, device="cuda")
    crp, cci, cv, brp, bci, bv = pkg.divide(0, 0, z, None, None, 16, 0.5)
    assert crp.cpu().tolist() == [0] and brp.cpu().tolist() == [0] and cci.numel() == 0 and bci.numel() == 0
    rp = torch.zeros(41, dtype=torch.int32, device="cuda")  # 40 empty rows
    crp, cci, cv, brp, bci, bv = pkg.divide(40, 40, rp, None, None, 16, 0.5)
    assert crp.cpu().tolist() == [0] * 41 and brp.cpu().tolist() == [0] * 4


@pytest.mark.parametrize("math,tol", [(0, TOL_FP32), (1, TOL_TC), (2, TOL_TC)])
def test_divide_hybrid_spmm_matches_csr(pkg, oracle, math, tol):
    """divide.cu:348-373: C = A_csr * B (beta = 1 on zeroed C), then C += A_bsr * B; the sum is A * B."""
    rng = np.random.default_rng(11)
    bs, K = 16, 128
    n = 64 * bs
    rp, ci = _banded_graph(rng, n, band=bs, p_dense=0.8, avg=6)
    v = rng.uniform(-1, 1, ci.size).astype(np.float32)
    B = rng.uniform(-1, 1, (n, K)).astype(np.float32)
    crp, cci, cv, brp, bci, bv = pkg.divide(n, n, dev(rp), dev(ci), dev(v), bs, 0.3)
    assert 0 < bci.numel() and 0 < cci.numel() < ci.size
    Bd = dev(B)
    C = torch.zeros(n, K, device="cuda")
    pkg.csrmm(crp, cci, cv, Bd, C, n, n, K, alpha=1.0, beta=1.0)
    mb = n // bs
    pkg.bsrmm(pkg.DIR_ROW, 1, mb, K, mb, bci.numel(), 1.0, bv, brp, bci, bs, Bd, K, 1.0, C, K,
              c_layout=pkg.ROW_MAJOR, math=math)
    ref = oracle.csrmm(rp, ci, v, B, K)
    assert err_vs(oracle, C.cpu().numpy(), ref, rp, ci, v, B, K) <= tol


def test_divide_full_size_proteins_property(pkg):
    """Size-independent property at the C3 (proteins-shaped) size: the two parts partition the non-zeros
    (counts add up, every block meets the threshold, no remainder entry lies in a dense block)."""
    from spmm_denseblock_b200 import graphs
    rp, ci = graphs.make_config_graph("c3_proteins", device="cuda", scale=0.25)
    n, nnz, bs, dens = rp.numel() - 1, ci.numel(), 16, 0.2
    crp, cci, cv, brp, bci, bv = pkg.divide(n, n, rp, ci, None, bs, dens)
    per_block = (bv.view(-1, bs * bs) != 0).sum(1)
    assert int(per_block.sum()) + cci.numel() == nnz
    assert int(per_block.min()) >= pkg.divide_threshold(bs, dens)
    # remainder entries never fall in a dense block of their block row
    mb = (n + bs - 1) // bs
    deg = (crp[1:] - crp[:-1]).long()
    rows = torch.arange(n, device="cuda").repeat_interleave(deg)
    rem_key = (rows // bs) * mb + (cci.long() // bs)
    bdeg = (brp[1:] - brp[:-1]).long()
    blk_key = torch.arange(mb, device="cuda").repeat_interleave(bdeg) * mb + bci.long()
    assert not torch.isin(rem_key, blk_key).any()


# ----------------------------------------------------------------------------- permutation on the device
def test_permute_golden_reference(pkg, golden_dir):
    """GPU permutation against the output of the reference's OWN permutate (reorder_strategy.cc:26-39,
    fixture made by tests/golden/make_golden.py): bit-exact."""
    import os
    d = np.load(os.path.join(golden_dir, "ref_permutate.npz"))
    nrp, nci, nv = pkg.permute_csr(dev(d["rowptr"]), dev(d["colind"]), None, dev(d["old2new"]))
    assert nv is None
    assert np.array_equal(nrp.cpu().numpy(), d["p_rowptr"]) and np.array_equal(nci.cpu().numpy(), d["p_colind"])


@pytest.mark.parametrize("key64", [0, 1])
@pytest.mark.parametrize("n,avg,hubs", [(700, 5, {}), (3000, 20, {5: 40, 9: 1000, 11: 1025, 77: 2999}),
                                        (9000, 12, {2: 32, 3: 33, 10: 128, 11: 129, 20: 512, 21: 513, 30: 4096,
                                                    31: 4097, 40: 2048, 41: 8192, 42: 8999, 43: 1024}),
                                        (40000, 3, {123: 16384, 9000: 16385, 39999: 33000})])
def test_permute_csr_bit_exact_with_values(pkg, oracle, n, avg, hubs, key64):
    """All row-length classes (warp with 1 / 4 / 16 keys per lane, shared-memory CTA small / large, global staging)
    and both key widths; values follow entries."""
    rng = np.random.default_rng(n)
    rp, ci, v = random_csr(rng, n, n, avg, hubs=hubs, empty_rows={1, n - 1})
    perm = rng.permutation(n).astype(np.int32)
    pkg.set_tuning("PERM_KEY64", key64)
    try:
        nrp, nci, nv = pkg.permute_csr(dev(rp), dev(ci), dev(v), dev(perm))
    finally:
        pkg.set_tuning("PERM_KEY64", None)
    orp, oci = oracle.permutate_csr(rp, ci, perm)
    assert np.array_equal(nrp.cpu().numpy(), orp) and np.array_equal(nci.cpu().numpy(), oci)
    # values: A'[perm[i], perm[j]] == A[i, j]
    import scipy.sparse as sp
    A = sp.csr_matrix((v, ci, rp), shape=(n, n))
    Ap = sp.csr_matrix((nv.cpu().numpy(), nci.cpu().numpy(), nrp.cpu().numpy()), shape=(n, n))
    inv = np.empty(n, np.int64)
    inv[perm] = np.arange(n)
    back = Ap[perm][:, perm].tocsr()
    back.sort_indices()
    assert np.array_equal(back.indptr, rp) and np.array_equal(back.indices, ci) and np.array_equal(back.data, v)


def test_permute_rows_heatmap_metrics_and_spmm_invariance(pkg, oracle):
    """P (A B) = (P A P^T)(P B): SpMM on the permuted operands, un-permuted, equals the original product;
    heat map and block metrics equal the reference definitions (utility.cc:47-88, reorder_graph.cc:12-24)."""
    rng = np.random.default_rng(4)
    n, K, bs = 2500, 64, 16
    rp, ci, v = random_csr(rng, n, n, 7, hubs={3: 900})
    perm = rng.permutation(n).astype(np.int32)
    inv = np.empty(n, np.int32)
    inv[perm] = np.arange(n, dtype=np.int32)
    B = rng.uniform(-1, 1, (n, K)).astype(np.float32)
    nrp, nci, nv = pkg.permute_csr(dev(rp), dev(ci), dev(v), dev(perm))
    Bp = pkg.permute_rows(dev(B), dev(perm))
    assert np.array_equal(Bp.cpu().numpy()[perm], B)
    Cp = torch.empty(n, K, device="cuda")
    pkg.csrmm(nrp, nci, nv, Bp, Cp, n, n, K)
    C = pkg.permute_rows(Cp, dev(inv)).cpu().numpy()
    ref = oracle.csrmm(rp, ci, v, B, K)
    assert err_vs(oracle, C, ref, rp, ci, v, B, K) <= TOL_FP32
    heat = pkg.heatmap(dev(rp), dev(ci), 256).cpu().numpy()
    nb = (n + 255) // 256
    want = np.zeros((nb, nb), np.int64)
    rows = np.repeat(np.arange(n), np.diff(rp))
    np.add.at(want, (rows // 256, ci // 256), 1)
    assert np.array_equal(heat, want)
    nnzb, density, util, average = pkg.block_metrics(dev(rp), dev(ci), bs)
    assert nnzb == oracle.calculate_nnzb(rp, ci, bs)
    nbb = (n + bs - 1) // bs
    assert density == nnzb / (nbb * nbb) and util == ci.size / (nnzb * bs * bs) and average == ci.size / nnzb


def test_permute_full_size_products_property(pkg):
    """Size-independent properties at the C5 (products-shaped) size: rows sorted, degrees moved, the inverse
    permutation restores the matrix bit for bit."""
    from spmm_denseblock_b200 import graphs
    rp, ci = graphs.make_config_graph("c5_products", device="cuda", scale=0.5)
    n = rp.numel() - 1
    g = torch.Generator(device="cuda").manual_seed(5)
    perm = torch.randperm(n, generator=g, device="cuda").to(torch.int32)
    inv = torch.empty_like(perm)
    inv[perm.long()] = torch.arange(n, device="cuda", dtype=torch.int32)
    nrp, nci, _ = pkg.permute_csr(rp, ci, None, perm)
    deg, ndeg = rp[1:] - rp[:-1], nrp[1:] - nrp[:-1]
    assert torch.equal(ndeg[perm.long()], deg)
    same_row = torch.ones(nci.numel() - 1, dtype=torch.bool, device="cuda")
    same_row[(nrp[1:-1].long() - 1).clamp_(min=0)] = False  # last entry of a row vs first of the next
    assert bool(((nci[1:] > nci[:-1]) | ~same_row).all())
    brp, bci, _ = pkg.permute_csr(nrp, nci, None, inv)
    assert torch.equal(brp, rp) and torch.equal(bci, ci)


# ----------------------------------------------------------------------------- orderings produced on the device
def _order_cases():
    from helpers import undirected_graph
    rng = np.random.default_rng(21)
    yield "components+isolated+hub+loops+duplicates", undirected_graph(
        rng, 5000, 4, components=9, isolated=(0, 17, 4999), hubs={3: 1500, 40: 300}, self_loops=20, duplicates=30)
    yield "unsorted adjacency lists", undirected_graph(rng, 3000, 5, components=2, shuffle_rows=True)
    yield "all isolated", (np.zeros(258, np.int32), np.zeros(0, np.int32))
    path = np.arange(1500)
    rp = np.zeros(1501, np.int32)
    rp[1:] = np.cumsum(np.where((path == 0) | (path == 1499), 1, 2))
    ci = np.concatenate([[1]] + [[i - 1, i + 1] for i in range(1, 1499)] + [[1498]]).astype(np.int32)
    yield "path (one level per vertex)", (rp, ci)
    yield "many small components", undirected_graph(rng, 20000, 2.2, components=4000)


@pytest.mark.parametrize("name,graph", list(_order_cases()), ids=lambda x: x if isinstance(x, str) else "")
def test_order_bfs_is_the_reference_permutation(pkg, oracle, name, graph):
    """spmmdb_order_bfs == BFSTraversal (reorder_strategy.cc:68-98) on undirected graphs: the permutation equals the
    sequential restatement's, and the graph permuted with it on the device equals the reference's own output graph."""
    rp, ci = graph
    p = pkg.order_bfs(dev(rp), dev(ci))
    assert np.array_equal(p.cpu().numpy(), oracle.bfs_order(rp, ci))
    if ci.size and oracle.ref_host_lib() is not None and hasattr(oracle.ref_host_lib(), "ref_bfs_graph"):
        nrp, nci, _ = pkg.permute_csr(dev(rp), dev(ci), None, p)
        wrp, wci = oracle.ref_bfs_graph(rp, ci)
        assert np.array_equal(nrp.cpu().numpy(), wrp) and np.array_equal(nci.cpu().numpy(), wci)


def test_order_bfs_config_shapes_and_directed_patterns(pkg, oracle):
    """The C2-shaped graph (169 343 vertices) and the C4 batch of molecule graphs (tens of thousands of components),
    symmetrised and as generated (directed), and small directed patterns where most vertices are reached from a
    smaller root's territory: the sequential restatement's permutation every time, and the reference's own output
    graph on the small ones."""
    import scipy.sparse as sp
    from spmm_denseblock_b200 import graphs
    for cfg, scale in (("c2_arxiv", 1.0), ("c4_molhiv", 0.25)):
        rp, ci = graphs.make_config_graph(cfg, device="cuda", scale=scale)
        p = pkg.order_bfs(rp, ci).cpu().numpy()
        rp, ci = rp.cpu().numpy(), ci.cpu().numpy()
        assert np.array_equal(p, oracle.bfs_order(rp, ci)), cfg + " as generated"
        n = rp.size - 1
        A = sp.csr_matrix((np.ones(ci.size, np.int8), ci, rp), shape=(n, n))
        A = ((A + A.T) > 0).astype(np.int8).tocsr()
        A.sort_indices()
        rp, ci = A.indptr.astype(np.int32), A.indices.astype(np.int32)
        p = pkg.order_bfs(dev(rp), dev(ci)).cpu().numpy()
        assert np.array_equal(p, oracle.bfs_order(rp, ci)), cfg + " symmetrised"
    rng = np.random.default_rng(77)
    cases = [random_csr(rng, 3000, 3000, 1.2)[:2],                       # sparse, directed: many roots, long chains
             random_csr(rng, 2000, 2000, 4, hubs={1999: 1500}, empty_rows={0, 1, 2})[:2],
             (np.array([0, 0, 1, 1], np.int32), np.array([0], np.int32)),  # 1 -> 0 only
             (np.array([0, 1, 2, 3, 3], np.int32), np.array([3, 0, 1], np.int32))]  # 0 -> 3, 1 -> 0, 2 -> 1
    has_ref = oracle.ref_host_lib() is not None and hasattr(oracle.ref_host_lib(), "ref_bfs_graph")
    for rp, ci in cases:
        p = pkg.order_bfs(dev(rp), dev(ci))
        assert np.array_equal(p.cpu().numpy(), oracle.bfs_order(rp, ci))
        if has_ref:
            nrp, nci, _ = pkg.permute_csr(dev(rp), dev(ci), None, p)
            wrp, wci = oracle.ref_bfs_graph(rp, ci)
            assert np.array_equal(nrp.cpu().numpy(), wrp) and np.array_equal(nci.cpu().numpy(), wci)


def test_orderings_golden_reference(pkg, golden_dir):
    """The device orderings against the committed outputs of the reference's own functions
    (tests/golden/ref_orderings.npz): the graph permuted with spmmdb_order_bfs equals BFSTraversal's output on
    undirected, unsorted-list and directed inputs; spmmdb_order_rcm equals reverseCuthillMcKee's where the
    reference's sort is stable; spmmdb_order_by_degree gives maxDegreeSort's row lengths."""
    import os
    d = np.load(os.path.join(golden_dir, "ref_orderings.npz"))
    for name in ("und", "unsorted", "directed", "lowdeg"):
        rp, ci = dev(d[name + "_rowptr"]), dev(d[name + "_colind"])
        nrp, nci, _ = pkg.permute_csr(rp, ci, None, pkg.order_bfs(rp, ci))
        assert np.array_equal(nrp.cpu().numpy(), d[name + "_bfs_rowptr"]), name
        assert np.array_equal(nci.cpu().numpy(), d[name + "_bfs_colind"]), name
        nrp, _, _ = pkg.permute_csr(rp, ci, None, pkg.order_by_degree(rp))
        assert np.array_equal(nrp.cpu().numpy(), d[name + "_deg_rowptr"]), name
        if name == "lowdeg":
            nrp, nci, _ = pkg.permute_csr(rp, ci, None, pkg.order_rcm(rp, ci))
            assert np.array_equal(nrp.cpu().numpy(), d[name + "_rcm_rowptr"])
            assert np.array_equal(nci.cpu().numpy(), d[name + "_rcm_colind"])


def test_order_rcm(pkg, oracle):
    """spmmdb_order_rcm == reverseCuthillMcKee (reorder_strategy.cc:57-66) with equal-degree neighbours in ascending
    id: the restatement's permutation on graphs with hubs and unsorted lists; the reference's own output graph where
    its std::sort is stable (maximum degree <= 16)."""
    from helpers import undirected_graph
    rng = np.random.default_rng(31)
    big = undirected_graph(rng, 6000, 6, components=3, isolated=(1, 2), hubs={4: 700, 90: 5000, 91: 40}, self_loops=10)
    low = undirected_graph(rng, 5000, 3, components=6, isolated=(0,))
    assert np.diff(low[0]).max() <= 16
    directed = random_csr(rng, 3000, 3000, 3, hubs={7: 900})[:2]
    for rp, ci in (big, low, directed, (np.zeros(4, np.int32), np.zeros(0, np.int32))):
        # ascending lists: the restatement's stable sort then breaks degree ties by ascending id, like the device
        p = pkg.order_rcm(dev(rp), dev(ci))
        assert np.array_equal(p.cpu().numpy(), oracle.rcm_order(rp, ci))
    rp, ci = low
    p = pkg.order_rcm(dev(rp), dev(ci))
    nrp, nci, _ = pkg.permute_csr(dev(rp), dev(ci), None, p)
    wrp, wci = oracle.ref_rcmk_graph(rp, ci)
    assert np.array_equal(nrp.cpu().numpy(), wrp) and np.array_equal(nci.cpu().numpy(), wci)


def test_order_by_degree(pkg, oracle):
    """spmmdb_order_by_degree == maxDegreeSort (reorder_strategy.cc:41-55) with ties in ascending id: the
    restatement's permutation exactly; the reference's row lengths always, its rows when all degrees differ."""
    rng = np.random.default_rng(8)
    rp, ci, _ = random_csr(rng, 30000, 30000, 6, hubs={5: 9000, 77: 20000, 29999: 513}, empty_rows={0, 9, 29998})
    q = pkg.order_by_degree(dev(rp))
    assert np.array_equal(q.cpu().numpy(), oracle.degree_order(rp))
    has_ref = oracle.ref_host_lib() is not None and hasattr(oracle.ref_host_lib(), "ref_max_degree_sort_graph")
    if has_ref:
        nrp, _, _ = pkg.permute_csr(dev(rp), dev(ci), None, q)
        assert np.array_equal(nrp.cpu().numpy(), oracle.ref_max_degree_sort_graph(rp, ci)[0])
    n = 700
    degs = rng.permutation(n)
    rows = [np.sort(rng.choice(n, size=d, replace=False)) for d in degs]
    rp = np.zeros(n + 1, np.int32)
    rp[1:] = np.cumsum(degs)
    ci = np.concatenate(rows).astype(np.int32)
    q = pkg.order_by_degree(dev(rp))
    assert np.array_equal(q.cpu().numpy(), oracle.degree_order(rp))
    if has_ref:
        nrp, nci, _ = pkg.permute_csr(dev(rp), dev(ci), None, q)
        wrp, wci = oracle.ref_max_degree_sort_graph(rp, ci)
        assert np.array_equal(nrp.cpu().numpy(), wrp) and np.array_equal(nci.cpu().numpy(), wci)
    assert pkg.order_by_degree(dev(np.zeros(5, np.int32))).cpu().tolist() == [0, 1, 2, 3]


@pytest.mark.parametrize("bs", [2, 3, 4, 8, 16, 32, 64])
def test_block_fill_histogram_divide_metrics_bit_exact(pkg, oracle, bs):
    """spmmdb_block_fill_histogram == calculate_block_density_dist (block_density_dist.cpp:47-86), every decile,
    on a matrix with sparse background and dense diagonal bands (so the upper deciles are populated)."""
    rng = np.random.default_rng(bs)
    n = 1000
    rp, ci = _banded_graph(rng, n, band=max(bs, 4), p_dense=0.85, avg=7)
    want = oracle.block_density_dist(rp, ci, bs)
    got = pkg.block_fill_histogram(dev(rp), dev(ci), bs)
    assert got == want.tolist(), (got, want.tolist())
    assert sum(got) == pkg.block_metrics(dev(rp), dev(ci), bs)[0]


@pytest.mark.parametrize("n,bs", [(1000, 16), (1000, 32), (1024, 32), (777, 16)])
@pytest.mark.parametrize("math,tol", [(0, TOL_FP32), (1, TOL_TC), (2, TOL_TC)])
def test_hybrid_divide_object_ragged_sizes(pkg, oracle, n, bs, math, tol):
    """spmmdb_hybrid_*: the split kept as one object, C = alpha * A * B + beta * C through CSR (remainder) + BSR
    (dense blocks), with m, n NOT multiples of the block size and neither B nor C padded (the BSR kernels clip).
    Guard rows after C catch a write past row m."""
    rng = np.random.default_rng(n + bs + math)
    K = 128
    rp, ci = _banded_graph(rng, n, band=bs, p_dense=0.8, avg=6)
    v = rng.uniform(-1, 1, ci.size).astype(np.float32)
    B = rng.uniform(-1, 1, (n, K)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (n, K)).astype(np.float32)
    h = pkg.Hybrid(n, n, dev(rp), dev(ci), dev(v), bs, 0.3)
    assert h.nnzb > 0 and 0 < h.csr_nnz < ci.size
    buf = torch.full((n + 64, K), 7.0, device="cuda")
    buf[:n] = dev(C0)
    Bbuf = torch.full((n + 64, K), float("nan"), device="cuda")  # rows past n must never be read into the result
    Bbuf[:n] = dev(B)
    h.spmm(Bbuf[:n], buf[:n], K, alpha=0.5, beta=-1.5, math=math)
    torch.cuda.synchronize()
    assert torch.all(buf[n:] == 7.0)
    ref = oracle.csrmm(rp, ci, v, B, K, C0=C0, alpha=0.5, beta=-1.5)
    scale = 0.5 * oracle.csrmm(rp, ci, np.abs(v), np.abs(B), K) + 1.5 * np.abs(C0)
    err = float(np.max(np.abs(buf[:n].cpu().numpy().astype(np.float64) - ref) / np.maximum(scale, 1e-30)))
    assert err <= tol, err
    h.close()

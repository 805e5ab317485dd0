"""spmmdb_csrmm_noalloc_f32 (csrc/noalloc.cu): the CSR x dense entry point whose scratch is a fixed __device__ array
inside the library, checked through the C ABI against the existing entry point spmmdb_csrmm_f32 on calls in a row.
Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-5
NOT_SUPPORTED = -2
MAX_K = 512  # SPMMDB_NOALLOC_MAX_K (include/spmmdb.h)


@pytest.fixture(scope="module")
def pkg():
    import spmm_denseblock_b200 as p
    assert torch.cuda.is_available()
    p.lib()
    return p


def make_csr(seed, m, k, avg, long_rows):
    """Random CSR with distinct sorted columns per row; long_rows: {row: length}; rows 1 and m - 1 are empty."""
    rng = np.random.default_rng(seed)
    deg = rng.poisson(avg, m).astype(np.int64)
    for r, d in long_rows.items():
        deg[r] = d
    deg[1] = deg[m - 1] = 0
    deg = np.minimum(deg, k)
    rp = np.zeros(m + 1, np.int32)
    rp[1:] = np.cumsum(deg)
    ci = np.empty(int(rp[-1]), np.int32)
    for r in range(m):
        d = int(deg[r])
        if d:
            ci[rp[r]:rp[r + 1]] = np.sort(rng.choice(k, d, replace=False)) if d > 64 else \
                np.unique(np.concatenate([rng.integers(0, k, d), np.arange(d)]))[:d]
    v = rng.uniform(-1, 1, ci.size).astype(np.float32)
    return rp, ci, v


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def p(t):
    return C.c_void_p(t.data_ptr())


def call_noalloc(pkg, rp, ci, v, B, ldb, Cout, ldc, m, k, K, alpha, beta):
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return pkg.lib().spmmdb_csrmm_noalloc_f32(m, k, K, p(rp), p(ci), p(v), p(B), ldb, p(Cout), ldc, alpha, beta, st)


def compare(pkg, rp, ci, v, m, k, K, alpha=1.0, beta=0.0, pad=0, seed=0):
    """Runs the new and the existing entry point on the same operands; returns the new result after asserting the
    relative error bound max |new - old| / (|alpha| |A| |B| + |beta| |C0|) <= 1e-5."""
    g = torch.Generator(device="cuda").manual_seed(1000 + seed)
    ld = K + pad
    B = torch.rand(k, ld, generator=g, device="cuda") * 2 - 1
    C0 = torch.rand(m, ld, generator=g, device="cuda") * 2 - 1
    got, want, scale = C0.clone(), C0.clone(), C0.abs()
    rc = call_noalloc(pkg, rp, ci, v, B, ld, got, ld, m, k, K, alpha, beta)
    assert rc == 0, pkg.lib().spmmdb_last_error()
    pkg.csrmm(rp, ci, v, B, want, m, k, K, ldb=ld, ldc=ld, alpha=alpha, beta=beta)
    pkg.csrmm(rp, ci, v.abs(), B.abs(), scale, m, k, K, ldb=ld, ldc=ld, alpha=abs(alpha), beta=abs(beta))
    torch.cuda.synchronize()
    err = ((got[:, :K] - want[:, :K]).abs() / scale[:, :K].clamp_min(1e-30)).max().item()
    assert err <= TOL_FP32, err
    if pad:  # columns beyond K are not touched
        assert torch.equal(got[:, K:], C0[:, K:])
    if beta == 0.0:
        deg = rp[1:] - rp[:-1]
        assert bool((got[:, :K][deg == 0] == 0).all())  # empty rows are written
    return got


@pytest.mark.parametrize("K", [64, 100, 256, 512])
def test_noalloc_matches_csrmm_on_calls_in_a_row(pkg, K):
    m = k = 3000
    longs = {5: 1024, 9: 1025, 11: 1500, 77: 2999, 500: 2048, 2998: 1537}
    rp1, ci1, v1 = (dev(a) for a in make_csr(K, m, k, 6, longs))
    rp2, ci2, v2 = (dev(a) for a in make_csr(K + 1, 1777, k, 11, {3: 1100}))
    first = compare(pkg, rp1, ci1, v1, m, k, K, seed=1)
    compare(pkg, rp2, ci2, v2, 1777, k, K, alpha=0.5, beta=-2.0, pad=4 if K % 4 == 0 else 3, seed=2)
    again = compare(pkg, rp1, ci1, v1, m, k, K, seed=1)  # same operands after another call used the array
    assert torch.equal(first, again)
    compare(pkg, rp1, ci1, v1, m, k, K, alpha=-1.25, beta=0.75, seed=3)


def test_noalloc_scratch_capacity_overflow(pkg):
    """More long-row segments than the fixed array holds at K = 512 (4096): the rest are summed in place, same sums."""
    K, m, k = 512, 1800, 4000
    longs = {r: 1536 + (r % 7) for r in range(100, 1700)}  # 1600 rows x 4 segments (the last one short) > 4096
    rp, ci, v = (dev(a) for a in make_csr(7, m, k, 3, longs))
    a = compare(pkg, rp, ci, v, m, k, K, seed=4)
    b = compare(pkg, rp, ci, v, m, k, K, seed=4)
    assert torch.equal(a, b)  # which rows overflowed may differ between calls; the sums do not


def test_noalloc_rejects_sizes_beyond_the_limit_and_allocates_nothing(pkg):
    m = k = 2000
    rp, ci, v = (dev(a) for a in make_csr(3, m, k, 8, {4: 1300}))
    L = pkg.lib()
    B = torch.zeros(k, MAX_K + 4, device="cuda")
    Cm = torch.zeros(m, MAX_K + 4, device="cuda")
    rc = call_noalloc(pkg, rp, ci, v, B, MAX_K + 4, Cm, MAX_K + 4, m, k, MAX_K + 4, 1.0, 0.0)
    assert rc == NOT_SUPPORTED
    msg = L.spmmdb_last_error().decode()
    assert "SPMMDB_NOALLOC_MAX_K" in msg and str(MAX_K) in msg
    assert call_noalloc(pkg, rp, ci, v, B, 8, Cm, MAX_K, m, k, 64, 1.0, 0.0) == -1  # ldb < K
    # no device memory is taken by the calls (the first call loads the kernels: measure after it)
    assert call_noalloc(pkg, rp, ci, v, B, MAX_K + 4, Cm, MAX_K + 4, m, k, 256, 1.0, 0.0) == 0
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for K in (32, 256, 512):
        assert call_noalloc(pkg, rp, ci, v, B, MAX_K + 4, Cm, MAX_K + 4, m, k, K, 1.0, 0.0) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0 - (1 << 20)  # a cudaMalloc would take at least 2 MB

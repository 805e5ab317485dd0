"""Parity of the kernels the benchmark times.

`csrmm_stream_kernel<VEC, U, MINB>` with the inline segment queue (semaphore + tickets + partial sums added
by the last finisher) is what BENCH / SCALE measure on the products-shaped graph; it is taken only for
nnz >= 4 M.  These tests force it on small matrices (tuning knobs CSR_SMALL = 0, CSR_INLINE = 1, segment 256)
for every VEC, and run it at its natural size on config-shaped inputs; every result is compared with the fp64
oracle (<= 1e-5 relative to sum |a||b|, gespmm_csrmm.h:95-166 semantics), bit for bit with the reference
kernel's sequential fp32 FMA chain on rows of at most one segment, twice (determinism) and against row shards.
"""
import numpy as np
import pytest
import torch

from helpers import random_csr

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-5
TOL_TC = 1e-2
TOL_BF16B = 1e-2  # CSR with bf16-stored B (fp32 accumulate): B is rounded to 8 mantissa bits


@pytest.fixture(scope="module")
def pkg():
    import spmm_denseblock_b200 as p
    assert torch.cuda.is_available()
    p.lib()
    return p


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def err_vs(oracle, got, ref64, rp, ci, v, Bm, K, alpha=1.0, beta=0.0, C0=None):
    scale = abs(alpha) * oracle.csrmm(rp, ci, np.abs(v), np.abs(Bm), K)
    if C0 is not None:
        scale = scale + abs(beta) * np.abs(C0)
    return float(np.max(np.abs(got.astype(np.float64) - ref64) / np.maximum(scale, 1e-30)))


@pytest.fixture(params=["registers", "bulk1", "bulk2"])
def forced_stream(pkg, request):
    """Small matrices through the benchmark's kernels: no small-matrix path, segments of 256 taken inline.
    `registers` = csrmm_stream_kernel, `bulk1` / `bulk2` = csrmm_lean_kernel at two occupancies (the same gather with
    the inner loop rebuilt around 8-aligned batches and a row-end mask)."""
    pkg.set_tuning("CSR_SMALL", 0)
    pkg.set_tuning("CSR_INLINE", 1)
    pkg.set_tuning("CSR_LEAN", {"registers": 0, "bulk1": 1, "bulk2": 2}[request.param])
    pkg.set_csrmm_segment(256)
    yield 256
    pkg.set_csrmm_segment(0)
    pkg.set_tuning("CSR_SMALL", None)
    pkg.set_tuning("CSR_INLINE", None)
    pkg.set_tuning("CSR_LEAN", None)


@pytest.fixture(params=[0, 1, 2])  # 0: csrmm_stream_kernel, 1 / 2: csrmm_lean_kernel (the default is 1)
def bulk_mode(pkg, request):
    pkg.set_tuning("CSR_LEAN", request.param)
    yield request.param
    pkg.set_tuning("CSR_LEAN", None)


def hub_matrix(seed, m=6000, k=9000, avg=8):
    rng = np.random.default_rng(seed)
    hubs = {0: 257, 1: 256, 17: 5000, 18: 3000, 19: 700, 1000: 8999, 2500: 513, 2501: 512, 2502: 511,
            m - 1: 4000, m - 2: 300}
    for r in rng.choice(np.arange(m // 2, m - 10), size=25, replace=False):
        hubs[int(r)] = int(rng.integers(257, 2500))
    return rng, random_csr(rng, m, k, avg, hubs=hubs, empty_rows={5, 6, 7, 2503, m - 3})


@pytest.mark.parametrize("K", [32, 64, 128, 256, 384, 512])
def test_stream_kernel_inline_segments_vs_oracle(pkg, oracle, forced_stream, K):
    rng, (rp, ci, v) = hub_matrix(K)
    m, k = rp.size - 1, 9000
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    C = torch.full((m, K), float("nan"), device="cuda")
    drp, dci, dv, dB = dev(rp), dev(ci), dev(v), dev(B)
    pkg.csrmm(drp, dci, dv, dB, C, m, k, K)
    got = C.cpu().numpy()
    ref = oracle.csrmm(rp, ci, v, B, K)
    assert err_vs(oracle, got, ref, rp, ci, v, B, K) <= TOL_FP32
    seq = oracle.csrmm_f32_seq(rp, ci, v, B, K)
    short = np.diff(rp) <= forced_stream
    assert np.array_equal(got[short], seq[short])
    assert np.all(got[[5, 6, 7, 2503, m - 3]] == 0)  # empty rows are written
    # determinism: who computes which segment changes from launch to launch, the bits must not
    for _ in range(3):
        C2 = torch.full((m, K), float("nan"), device="cuda")
        pkg.csrmm(drp, dci, dv, dB, C2, m, k, K)
        assert torch.equal(C, C2)
    # nnz unknown to the callee (gespmm_csrmm's signature)
    C3 = torch.full((m, K), float("nan"), device="cuda")
    pkg.gespmm_csrmm(m, K, drp, dci, dv, dB, C3)
    assert torch.equal(C, C3)


@pytest.mark.parametrize("K", [64, 256, 384, 512])
@pytest.mark.parametrize("alpha,beta", [(0.5, -2.0), (2.0, 1.0)])
def test_stream_kernel_inline_segments_alpha_beta(pkg, oracle, forced_stream, K, alpha, beta):
    rng, (rp, ci, v) = hub_matrix(100 + K)
    m, k = rp.size - 1, 9000
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (m, K)).astype(np.float32)
    C = dev(C0)
    pkg.csrmm(dev(rp), dev(ci), dev(v), dev(B), C, m, k, K, alpha=alpha, beta=beta)
    ref = oracle.csrmm(rp, ci, v, B, K, C0=C0, alpha=alpha, beta=beta)
    assert err_vs(oracle, C.cpu().numpy(), ref, rp, ci, v, B, K, alpha, beta, C0) <= TOL_FP32


@pytest.mark.parametrize("K", [128, 256, 512])
def test_stream_kernel_understated_nnz_void_tickets(pkg, oracle, forced_stream, K):
    """A caller that understates nnz gets a queue too small for its hub rows: rows that do not fit publish void
    tickets and are computed in place by the warp that met them."""
    rng, (rp, ci, v) = hub_matrix(200 + K)
    m, k = rp.size - 1, 9000
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    ref = oracle.csrmm(rp, ci, v, B, K)
    for nnz in (256, 3000, 20000):
        C = torch.full((m, K), float("nan"), device="cuda")
        pkg.csrmm(dev(rp), dev(ci), dev(v), dev(B), C, m, k, K, nnz=nnz)
        assert err_vs(oracle, C.cpu().numpy(), ref, rp, ci, v, B, K) <= TOL_FP32, nnz


@pytest.mark.parametrize("K", [128, 256])
def test_stream_kernel_shards_concatenate_bit_for_bit(pkg, oracle, forced_stream, K):
    """8 nnz-balanced row ranges, each run as its own call with rebased rowptr and the whole
    matrix's segment length, concatenate to the single-call result bit for bit."""
    from spmm_denseblock_b200 import shard
    rng, (rp, ci, v) = hub_matrix(300 + K)
    m, k = rp.size - 1, 9000
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    dB = dev(B)
    full = torch.empty(m, K, device="cuda")
    pkg.csrmm(dev(rp), dev(ci), dev(v), dB, full, m, k, K)
    ranges = shard.shard_plan(torch.from_numpy(rp), 8)
    assert ranges[0][0] == 0 and ranges[-1][1] == m
    for lo, hi in ranges:
        rps = (rp[lo:hi + 1] - rp[lo]).astype(np.int32)
        cis, vs = ci[rp[lo]:rp[hi]], v[rp[lo]:rp[hi]]
        part = torch.full((hi - lo, K), float("nan"), device="cuda")
        pkg.csrmm(dev(rps), dev(cis), dev(vs), dB, part, hi - lo, k, K)
        assert torch.equal(part, full[lo:hi]), (lo, hi)


def test_stream_kernel_two_streams_in_flight(pkg, oracle, forced_stream):
    """Library scratch is per (device, stream): two calls on two streams in flight together do not share the
    long-row queue."""
    K = 256
    rng, (rp, ci, v) = hub_matrix(7)
    rng2, (rp2, ci2, v2) = hub_matrix(8, m=5000, k=9000, avg=6)
    m, m2, k = rp.size - 1, rp2.size - 1, 9000
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    dB = dev(B)
    a = [dev(x) for x in (rp, ci, v)]
    b = [dev(x) for x in (rp2, ci2, v2)]
    ref1 = torch.empty(m, K, device="cuda")
    ref2 = torch.empty(m2, K, device="cuda")
    pkg.csrmm(*a, dB, ref1, m, k, K)
    pkg.csrmm(*b, dB, ref2, m2, k, K)
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    outs = []
    for it in range(6):
        c1 = torch.full((m, K), float("nan"), device="cuda")
        c2 = torch.full((m2, K), float("nan"), device="cuda")
        torch.cuda.synchronize()
        with torch.cuda.stream(s1):
            pkg.csrmm(*a, dB, c1, m, k, K)
        with torch.cuda.stream(s2):
            pkg.csrmm(*b, dB, c2, m2, k, K)
        outs.append((c1, c2))
    torch.cuda.synchronize()
    for c1, c2 in outs:
        assert torch.equal(c1, ref1) and torch.equal(c2, ref2)


# ----------------------------------------------------------------------------- config-shaped inputs, natural dispatch
def _config_case(name, scale, K, seed=1234, **over):
    from spmm_denseblock_b200 import graphs
    rp, ci = graphs.make_config_graph(name, device="cuda", scale=scale, seed=seed, **over)
    n = rp.numel() - 1
    g = torch.Generator(device="cuda").manual_seed(seed)
    B = torch.rand(n, K, generator=g, device="cuda") * 2 - 1
    val = torch.rand(ci.numel(), generator=g, device="cuda") * 2 - 1
    return rp, ci, val, B, n


@pytest.mark.parametrize("name,scale,K,over", [
    ("c2_arxiv", 1.0, 128, {}),                      # small-matrix path (pipe kernel + side stream)
    ("c4_molhiv", 1.0, 300, {}),                     # stream kernel VEC 3, 64-entry segments
    ("c5_products", 0.1, 256, {}),                   # stream kernel VEC 2 with inline segments at its natural size
    ("c5_products", 0.1, 256, dict(p_local=0.9, window=4096)),
    ("c5_products", 0.1, 512, {}),                   # VEC 4
    ("c5_products", 0.1, 128, {}),                   # VEC 1
    ("c3_proteins", 0.25, 128, {}),                  # rows kernel (271 nnz per row)
])
def test_csrmm_config_shapes_vs_oracle(pkg, oracle, bulk_mode, name, scale, K, over):
    if bulk_mode and name != "c5_products":
        pytest.skip("the lean kernel serves the stream kernel's shapes with inline segments (nnz >= 4 M)")
    rp, ci, val, B, n = _config_case(name, scale, K, **over)
    C = torch.full((n, K), float("nan"), device="cuda")
    pkg.csrmm(rp, ci, val, B, C, n, n, K)
    C2 = torch.full((n, K), float("nan"), device="cuda")
    pkg.gespmm_csrmm(n, K, rp, ci, val, B, C2)
    assert torch.equal(C, C2)
    h = [t.cpu().numpy() for t in (rp, ci, val, B)]
    ref = oracle.csrmm(*h, K)
    got = C.cpu().numpy()
    assert err_vs(oracle, got, ref, *h, K) <= TOL_FP32
    seq = oracle.csrmm_f32_seq(*h, K)
    short = np.diff(h[0]) <= pkg.csrmm_segment(ci.numel(), m=n)
    assert short.sum() > 0 and np.array_equal(got[short], seq[short])
    if name == "c5_products":
        assert ci.numel() >= 4 << 20 and (~short).sum() > 0  # inline-segment path with queued hub rows


def test_csrmm_column_major_c_beyond_65535_row_tiles(pkg, oracle):
    """cuSPARSE-layout call (column-major C, run_csrmm.cu:133-142) with more than 65 535 * 32 rows: the
    transposing epilogue must not put the row tiles on grid.y."""
    m, k, K = 2_200_000, 1000, 4
    rng = np.random.default_rng(5)
    rp = np.arange(m + 1, dtype=np.int32)  # one entry per row
    ci = rng.integers(0, k, m).astype(np.int32)
    v = rng.uniform(-1, 1, m).astype(np.float32)
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    C = torch.full((K, m), float("nan"), device="cuda")
    pkg.csrmm(dev(rp), dev(ci), dev(v), dev(B), C, m, k, K, c_layout=pkg.COL_MAJOR)
    ref = (v[:, None] * B[ci]).astype(np.float32)
    assert np.array_equal(C.cpu().numpy().T, ref)


def test_tuning_knobs_roundtrip(pkg):
    assert pkg.get_tuning("NO_SUCH_KNOB", 7) == 7
    pkg.set_tuning("NO_SUCH_KNOB", 3)
    assert pkg.get_tuning("NO_SUCH_KNOB", 7) == 3
    pkg.set_tuning("NO_SUCH_KNOB", None)
    assert pkg.get_tuning("NO_SUCH_KNOB", 7) == 7


# ----------------------------------------------------------------------------- extended entry: bf16 B, hot columns
def _bf16_round(x):
    return torch.from_numpy(x).to(torch.bfloat16).to(torch.float32).numpy()


@pytest.mark.parametrize("K", [64, 128, 256, 512, 1024])
@pytest.mark.parametrize("alpha,beta", [(1.0, 0.0), (0.5, -2.0)])
def test_csrmm_ex_bf16_b(pkg, oracle, bulk_mode, K, alpha, beta):
    """B stored as bf16, fp32 accumulate: exactly the fp32 kernel's arithmetic on the rounded B (bit for bit on
    rows of one segment), and within the stated 1e-2 of the product with the unrounded B."""
    rng, (rp, ci, v) = hub_matrix(400 + K, m=3000, k=5000)
    m, k = rp.size - 1, 5000
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (m, K)).astype(np.float32)
    Bb = pkg.dense_pack_bf16(dev(B))
    Br = _bf16_round(B)
    assert torch.equal(Bb.float().cpu(), torch.from_numpy(Br))  # round to nearest even, like torch
    C = dev(C0)
    pkg.csrmm_ex(dev(rp), dev(ci), dev(v), Bb, C, m, k, K, alpha=alpha, beta=beta)
    got = C.cpu().numpy()
    ref_r = oracle.csrmm(rp, ci, v, Br, K, C0=C0, alpha=alpha, beta=beta)
    assert err_vs(oracle, got, ref_r, rp, ci, v, Br, K, alpha, beta, C0) <= TOL_FP32
    ref = oracle.csrmm(rp, ci, v, B, K, C0=C0, alpha=alpha, beta=beta)
    assert err_vs(oracle, got, ref, rp, ci, v, B, K, alpha, beta, C0) <= TOL_BF16B
    if alpha == 1.0 and beta == 0.0:
        seq = oracle.csrmm_f32_seq(rp, ci, v, Br, K)
        short = np.diff(rp) <= pkg.csrmm_segment(ci.size, m=m)
        assert np.array_equal(got[short], seq[short])


def test_csrmm_ex_bf16_b_unsupported_shapes(pkg):
    z = torch.zeros(8, dtype=torch.int32, device="cuda")
    Bb = torch.zeros(4, 12, dtype=torch.bfloat16, device="cuda")
    C = torch.zeros(7, 12, device="cuda")
    with pytest.raises(pkg.SpmmdbError):
        pkg.csrmm_ex(z, z[:0], torch.zeros(0, device="cuda"), Bb, C, 7, 4, 12)  # K % 8 != 0


@pytest.mark.parametrize("K,bf16", [(256, False), (128, False), (256, True), (512, True)])
def test_csrmm_ex_hot_columns_same_bits(pkg, oracle, forced_stream, K, bf16):
    """The hot-column bitmap only changes L2 eviction hints: results are the bits of the plain call."""
    rng, (rp, ci, v) = hub_matrix(500 + K)
    m, k = rp.size - 1, 9000
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    dB = pkg.dense_pack_bf16(dev(B)) if bf16 else dev(B)
    d = [dev(x) for x in (rp, ci, v)]
    budget = 700
    hot, n_hot = pkg.csr_hot_columns(k, d[1], budget)
    cnt = np.bincount(ci, minlength=k)
    # whole frequency classes, as many as fit the budget
    t = max(1, min(x for x in range(1, cnt.max() + 2) if (cnt >= x).sum() <= budget))
    want = cnt >= t
    assert n_hot == int(want.sum()) <= budget and n_hot > 0
    bits = np.unpackbits(hot.cpu().numpy().view(np.uint8), bitorder="little")[:k].astype(bool)
    assert np.array_equal(bits, want)
    plain = torch.full((m, K), float("nan"), device="cuda")
    hinted = torch.full((m, K), float("nan"), device="cuda")
    pkg.csrmm_ex(*d, dB, plain, m, k, K)
    for policy in (0, 1, 2):
        pkg.set_tuning("CSR_COLD_POLICY", policy)
        try:
            pkg.csrmm_ex(*d, dB, hinted, m, k, K, hot=hot)
        finally:
            pkg.set_tuning("CSR_COLD_POLICY", None)
        assert torch.equal(plain, hinted), policy
    if not bf16:
        ref = oracle.csrmm(rp, ci, v, B, K)
        assert err_vs(oracle, plain.cpu().numpy(), ref, rp, ci, v, B, K) <= TOL_FP32


# ----------------------------------------------------------------------------- CUDA graph capture
@pytest.mark.parametrize("path", ["small_matrix", "stream_inline", "lean_inline"])
def test_csrmm_call_replays_correctly_from_a_cuda_graph(pkg, oracle, path):
    """A call (after one warm-up call on the capture stream, so that no scratch is allocated during capture; nnz
    passed, so that nothing synchronises) can be captured into a CUDA graph -- including the small-matrix path's side stream -- and every
    replay gets a fresh call number on the device: stale queue entries of the previous replay never look current."""
    if path != "small_matrix":
        pkg.set_tuning("CSR_SMALL", 0)
        pkg.set_tuning("CSR_INLINE", 1)
        pkg.set_tuning("CSR_LEAN", 1 if path == "lean_inline" else 0)
        pkg.set_csrmm_segment(256)
    try:
        K = 128
        rng, (rp, ci, v) = hub_matrix(77)
        m, k = rp.size - 1, 9000
        d = [dev(x) for x in (rp, ci, v)]
        Bs = [rng.uniform(-1, 1, (k, K)).astype(np.float32) for _ in range(3)]
        B = dev(Bs[0])
        C = torch.empty(m, K, device="cuda")
        cap = torch.cuda.Stream()
        with torch.cuda.stream(cap):  # warm-up ON THE CAPTURE STREAM: library scratch is per (device, stream)
            pkg.csrmm(*d, B, C, m, k, K, nnz=ci.size)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=cap):
            pkg.csrmm(*d, B, C, m, k, K, nnz=ci.size)
        for it in (1, 2, 0, 1):
            B.copy_(dev(Bs[it]))
            C.fill_(float("nan"))
            g.replay()
            torch.cuda.synchronize()
            ref = oracle.csrmm(rp, ci, v, Bs[it], K)
            assert err_vs(oracle, C.cpu().numpy(), ref, rp, ci, v, Bs[it], K) <= TOL_FP32, (path, it)
    finally:
        pkg.set_csrmm_segment(0)
        for knob in ("CSR_SMALL", "CSR_INLINE", "CSR_LEAN"):
            pkg.set_tuning(knob, None)


# ----------------------------------------------------------------------------- tile::gather4 variant (opt-in)
@pytest.mark.parametrize("K,bf16", [(256, False), (128, False), (64, False), (256, True), (128, True)])
@pytest.mark.parametrize("mode", [1, 2])
def test_tma_gather4_kernel_same_bits(pkg, oracle, K, bf16, mode):
    """csrmm_tma_kernel (CSR_TMA knob): the B-row gather as cp.async.bulk.tensor tile::gather4 copies into a
    shared-memory ring, the lean kernel's multiply loop -- same bits as the default kernel, alpha / beta included.
    Measured slower than the register gather (profiles/r2n), so nothing selects it by default."""
    rng, (rp, ci, v) = hub_matrix(600 + K)
    m, k = rp.size - 1, 9000
    B = rng.uniform(-1, 1, (k, K)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (m, K)).astype(np.float32)
    dB = pkg.dense_pack_bf16(dev(B)) if bf16 else dev(B)
    d = [dev(x) for x in (rp, ci, v)]
    pkg.set_tuning("CSR_SMALL", 0)
    pkg.set_tuning("CSR_INLINE", 1)
    pkg.set_csrmm_segment(256)
    try:
        for alpha, beta in ((1.0, 0.0), (0.5, -2.0)):
            ref, out = dev(C0), dev(C0)
            pkg.csrmm_ex(*d, dB, ref, m, k, K, alpha=alpha, beta=beta)
            pkg.set_tuning("CSR_TMA", mode)
            pkg.csrmm_ex(*d, dB, out, m, k, K, alpha=alpha, beta=beta)
            pkg.set_tuning("CSR_TMA", None)
            assert torch.equal(ref, out), (alpha, beta)
        if not bf16:
            want = oracle.csrmm(rp, ci, v, B, K, C0=C0, alpha=0.5, beta=-2.0)
            assert err_vs(oracle, out.cpu().numpy(), want, rp, ci, v, B, K, 0.5, -2.0, C0) <= TOL_FP32
    finally:
        pkg.set_tuning("CSR_TMA", None)
        pkg.set_csrmm_segment(0)
        pkg.set_tuning("CSR_SMALL", None)
        pkg.set_tuning("CSR_INLINE", None)

"""ctypes/numpy front-end of the CPU oracle (oracle/spmm_oracle.c) and of the
reference's own CPU code built into oracle/_ref (oracle/Makefile).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package
(spmm_denseblock_b200) never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")

ROW_MAJOR, COL_MAJOR = 0, 1


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference is present)."""
    so = os.path.join(HERE, "liboracle.so")
    src = os.path.join(HERE, "spmm_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", HERE, "all"])
    elif os.path.isdir("/root/reference") and not os.path.exists(os.path.join(HERE, "_ref", "libref_host.so")):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])
    return so


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.oracle_csr2bsr_nnz.restype = C.c_int64
        _lib.oracle_calculate_nnzb.restype = C.c_int64
        _lib.oracle_divide_matrix.restype = C.c_int64
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _opt(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def csrmm(rowptr, colind, val, B, K, ldb=None, b_layout=ROW_MAJOR, C0=None, ldc=None,
          c_layout=ROW_MAJOR, alpha=1.0, beta=0.0):
    """fp64-accumulated alpha*A*B + beta*C0 -> float64 [m, K] (row-major)."""
    rowptr, colind, val, B = _i32(rowptr), _i32(colind), _f32(val), _f32(B)
    m = rowptr.size - 1
    k = B.size // K if ldb is None else None
    if ldb is None:
        ldb = K if b_layout == ROW_MAJOR else k
    if ldc is None:
        ldc = K if c_layout == ROW_MAJOR else m
    out = np.zeros((m, K), np.float64)
    c0 = None if C0 is None else _f32(C0)
    f = lib().oracle_csrmm_f64acc
    f.argtypes = [C.c_int, C.c_int, _i32p, _i32p, _f32p, _f32p, C.c_int64, C.c_int, C.c_void_p,
                  C.c_int64, C.c_int, C.c_float, C.c_float, _f64p]
    f(m, K, rowptr, colind, val, B.ravel(), ldb, b_layout, _opt(c0), ldc, c_layout, alpha, beta, out)
    return out


def csrmm_f32_seq(rowptr, colind, val, B, K):
    """The reference kernel's own fp32 sequential-FMA arithmetic (row-major)."""
    rowptr, colind, val, B = _i32(rowptr), _i32(colind), _f32(val), _f32(B)
    m = rowptr.size - 1
    out = np.zeros((m, K), np.float32)
    f = lib().oracle_csrmm_f32_seq
    f.argtypes = [C.c_int, C.c_int, _i32p, _i32p, _f32p, _f32p, C.c_int64, _f32p, C.c_int64]
    f(m, K, rowptr, colind, val, B.ravel(), K, out, K)
    return out


def spmm_cc_csr(indptr, indices, dense, out=None):
    """spmm.cc:7-25 restated (double, int64, unit weights)."""
    indptr = np.ascontiguousarray(indptr, np.int64)
    indices = np.ascontiguousarray(indices, np.int64)
    dense = np.ascontiguousarray(dense, np.float64)
    if out is None:
        out = np.zeros((indptr.size - 1, dense.shape[1]), np.float64)
    f = lib().oracle_spmm_cc_csr
    f.argtypes = [C.c_int64, C.c_int64, _i64p, _i64p, _f64p, _f64p]
    f(indptr.size - 1, dense.shape[1], indptr, indices, dense, out)
    return out


def spmm_cc_csr_f32(rowptr, colind, val, B):
    rowptr, colind, val, B = _i32(rowptr), _i32(colind), _f32(val), _f32(B)
    out = np.zeros((rowptr.size - 1, B.shape[1]), np.float32)
    f = lib().oracle_spmm_cc_csr_f32
    f.argtypes = [C.c_int, C.c_int, _i32p, _i32p, _f32p, _f32p, _f32p]
    f(rowptr.size - 1, B.shape[1], rowptr, colind, val, B, out)
    return out


def bsrmm(bsr_rowptr, bsr_colind, bsr_val, bs, B, n, ldb, trans_b, direction=0, alpha=1.0,
          beta=0.0, C0=None, ldc=None, c_layout=COL_MAJOR, kb=0):
    """fp64-accumulated BSR SpMM -> float64 [mb*bs, n] (row-major)."""
    bsr_rowptr, bsr_colind, bsr_val, B = _i32(bsr_rowptr), _i32(bsr_colind), _f32(bsr_val), _f32(B)
    mb = bsr_rowptr.size - 1
    m = mb * bs
    if ldc is None:
        ldc = m if c_layout == COL_MAJOR else n
    out = np.zeros((m, n), np.float64)
    c0 = None if C0 is None else _f32(C0)
    f = lib().oracle_bsrmm_f64acc
    f.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, _f32p, _i32p,
                  _i32p, _f32p, C.c_int64, C.c_float, C.c_void_p, C.c_int64, C.c_int, _f64p]
    f(direction, int(trans_b), mb, n, kb, bs, alpha, bsr_val.ravel(), bsr_rowptr, bsr_colind,
      B.ravel(), ldb, beta, _opt(c0), ldc, c_layout, out)
    return out


def block_gemm(bsr_rowptr, bsr_colind, bsr_val, bs, Y, dim, Z0):
    """block_cublas.cu:123-136 -> float64 [m, dim] (row-major view of col-major z)."""
    bsr_rowptr, bsr_colind, bsr_val = _i32(bsr_rowptr), _i32(bsr_colind), _f32(bsr_val)
    Y, Z0 = _f32(Y), _f32(Z0)
    mb = bsr_rowptr.size - 1
    out = np.zeros((mb * bs, dim), np.float64)
    f = lib().oracle_block_gemm_f64acc
    f.argtypes = [C.c_int, C.c_int, C.c_int, _i32p, _i32p, _f32p, _f32p, _f32p, _f64p]
    f(mb, dim, bs, bsr_rowptr, bsr_colind, bsr_val.ravel(), Y.ravel(), Z0.ravel(), out)
    return out


def csr2bsr(rowptr, colind, val, m, n, bs, direction=0):
    """-> (bsr_rowptr[mb+1], bsr_colind[nnzb], bsr_val[nnzb, bs, bs]) bit-exact contract."""
    rowptr, colind, val = _i32(rowptr), _i32(colind), _f32(val)
    mb = (m + bs - 1) // bs
    brp = np.zeros(mb + 1, np.int32)
    f = lib().oracle_csr2bsr_nnz
    f.argtypes = [C.c_int, C.c_int, _i32p, _i32p, C.c_int, _i32p]
    nnzb = int(f(m, n, rowptr, colind, bs, brp))
    bci = np.zeros(max(nnzb, 1), np.int32)
    bv = np.zeros(max(nnzb, 1) * bs * bs, np.float32)
    g = lib().oracle_csr2bsr_fill
    g.argtypes = [C.c_int, C.c_int, C.c_int, _i32p, _i32p, _f32p, C.c_int, _i32p, _i32p, _f32p]
    g(direction, m, n, rowptr, colind, val, bs, brp, bci, bv)
    return brp, bci[:nnzb], bv[:nnzb * bs * bs].reshape(nnzb, bs, bs)


def bsr2csr(bsr_rowptr, bsr_colind, bsr_val, bs, direction=0):
    bsr_rowptr, bsr_colind, bsr_val = _i32(bsr_rowptr), _i32(bsr_colind), _f32(bsr_val)
    mb = bsr_rowptr.size - 1
    nnz = int(bsr_rowptr[mb]) * bs * bs
    rp = np.zeros(mb * bs + 1, np.int32)
    ci = np.zeros(max(nnz, 1), np.int32)
    v = np.zeros(max(nnz, 1), np.float32)
    f = lib().oracle_bsr2csr
    f.argtypes = [C.c_int, C.c_int, C.c_int, _i32p, _i32p, _f32p, C.c_int, _i32p, _i32p, _f32p]
    f(direction, mb, 0, bsr_rowptr, bsr_colind, bsr_val.ravel(), bs, rp, ci, v)
    return rp, ci[:nnz], v[:nnz]


def permutate_csr(rowptr, colind, old2new):
    rowptr, colind, old2new = _i32(rowptr), _i32(colind), _i32(old2new)
    n = rowptr.size - 1
    nrp = np.zeros(n + 1, np.int32)
    nci = np.zeros(max(colind.size, 1), np.int32)
    f = lib().oracle_permutate_csr
    f.argtypes = [C.c_int, _i32p, _i32p, _i32p, _i32p, _i32p]
    f(n, rowptr, colind, old2new, nrp, nci)
    return nrp, nci[:colind.size]


def bfs_order(rowptr, colind):
    """old2new of BFSTraversal (reorder_strategy.cc:68-98); colind in adjacency order."""
    rowptr, colind = _i32(rowptr), _i32(colind)
    n = rowptr.size - 1
    out = np.zeros(max(n, 1), np.int32)
    f = lib().oracle_bfs_order
    f.argtypes = [C.c_int, _i32p, _i32p, _i32p]
    f(n, rowptr, colind if colind.size else np.zeros(1, np.int32), out)
    return out[:n]


def rcm_order(rowptr, colind):
    """old2new of reverseCuthillMcKee (reorder_strategy.cc:57-66): every list sorted by descending degree of the
    neighbour (ties keep list order, i.e. a stable sort; the reference's std::sort is stable only for lists of at
    most 16 entries), then BFSTraversal."""
    rowptr, colind = _i32(rowptr), _i32(colind)
    n = rowptr.size - 1
    deg = np.diff(rowptr)
    rows = np.repeat(np.arange(n), deg)
    order = np.lexsort((np.arange(colind.size), -deg[colind].astype(np.int64), rows))
    return bfs_order(rowptr, colind[order])


def degree_order(rowptr):
    """old2new of maxDegreeSort (reorder_strategy.cc:41-55), equal degrees in ascending id."""
    rowptr = _i32(rowptr)
    n = rowptr.size - 1
    out = np.zeros(max(n, 1), np.int32)
    f = lib().oracle_degree_order
    f.argtypes = [C.c_int, _i32p, _i32p]
    f(n, rowptr, out)
    return out[:n]


def calculate_nnzb(rowptr, colind, bs):
    rowptr, colind = _i32(rowptr), _i32(colind)
    f = lib().oracle_calculate_nnzb
    f.argtypes = [C.c_int, _i32p, _i32p, C.c_int]
    return int(f(rowptr.size - 1, rowptr, colind, bs))


def divide_matrix(rowptr, colind, bs, density, val=None):
    """divide.cu:52-127 -> (csr_rowptr, csr_colind, bsr_rowptr, bsr_colind, bsr_val); with val (the
    reference uses unit weights) also csr_val as a sixth item."""
    rowptr, colind = _i32(rowptr), _i32(colind)
    n = rowptr.size - 1
    nb = (n + bs - 1) // bs
    f = lib().oracle_divide_matrix
    f.argtypes = [C.c_int, _i32p, _i32p, C.c_void_p, C.c_int, C.c_float] + [C.c_void_p] * 6 + [C.POINTER(C.c_int64)]
    v = None if val is None else _f32(val)
    vptr = None if v is None else v.ctypes.data
    cnnz = C.c_int64(0)
    nnzb = int(f(n, rowptr, colind, vptr, bs, density, None, None, None, None, None, None, C.byref(cnnz)))
    crp = np.zeros(n + 1, np.int32)
    cci = np.zeros(max(cnnz.value, 1), np.int32)
    cv = np.zeros(max(cnnz.value, 1), np.float32)
    brp = np.zeros(nb + 1, np.int32)
    bci = np.zeros(max(nnzb, 1), np.int32)
    bv = np.zeros(max(nnzb, 1) * bs * bs, np.float32)
    f(n, rowptr, colind, vptr, bs, density, crp.ctypes.data, cci.ctypes.data, cv.ctypes.data, brp.ctypes.data,
      bci.ctypes.data, bv.ctypes.data, C.byref(cnnz))
    out = (crp, cci[:cnnz.value], brp, bci[:nnzb], bv[:nnzb * bs * bs].reshape(nnzb, bs, bs))
    return out if val is None else out + (cv[:cnnz.value],)


# --------------------------------------------------------------------------
# The reference's own CPU code (oracle/_ref), when built.
# --------------------------------------------------------------------------
_ref_spmm = None
_ref_host = None


def ref_spmm_lib():
    """libref_spmm.so = /root/reference/spmm.cc + matrix.h shim, or None."""
    global _ref_spmm
    p = os.path.join(HERE, "_ref", "libref_spmm.so")
    if _ref_spmm is None and os.path.exists(p):
        _ref_spmm = C.CDLL(p)
    return _ref_spmm


def ref_host_lib():
    global _ref_host
    p = os.path.join(HERE, "_ref", "libref_host.so")
    if _ref_host is None and os.path.exists(p):
        try:
            _ref_host = C.CDLL(p)
        except OSError:
            _ref_host = None
    return _ref_host


def ref_csr_spmm(indptr, indices, dense, num_cols=None, out=None):
    """The reference's csr_spmm (spmm.cc:7-25) itself: double, int64, unit weights.  `out` (float64 [rows, K],
    fully overwritten) lets a timed caller keep the allocation outside its clock, as spmm.cc:63-72 does."""
    L = ref_spmm_lib()
    indptr = np.ascontiguousarray(indptr, np.int64)
    indices = np.ascontiguousarray(indices, np.int64)
    dense = np.ascontiguousarray(dense, np.float64)
    if out is None:
        out = np.zeros((indptr.size - 1, dense.shape[1]), np.float64)
    L.ref_csr_spmm.argtypes = [C.c_int64, C.c_int64, C.c_int64, _i64p, _i64p, _f64p, _f64p]
    L.ref_csr_spmm(indptr.size - 1, dense.shape[0] if num_cols is None else num_cols,
                   dense.shape[1], indptr, indices, dense, out)
    return out


def ref_omp_threads():
    L = ref_spmm_lib()
    return int(L.ref_omp_max_threads()) if L is not None else 0


def ref_set_omp_threads(n):
    """Give the reference's OpenMP loop n threads (torchrun exports OMP_NUM_THREADS=1 to its workers)."""
    L = ref_spmm_lib()
    if L is not None and hasattr(L, "ref_omp_set_threads"):
        L.ref_omp_set_threads(int(n))


def ref_calculate_nnzb(rowptr, colind, bs):
    L = ref_host_lib()
    rowptr, colind = _i32(rowptr), _i32(colind)
    L.ref_calculate_nnzb.argtypes = [C.c_int, _i32p, _i32p, C.c_int]
    return int(L.ref_calculate_nnzb(rowptr.size - 1, rowptr, colind, bs))


def block_density_dist(rowptr, colind, bs):
    """calculate_block_density_dist restated (block_density_dist.cpp:47-86): non-empty blocks per fill decile,
    decile = (int)((float)(count * 1.0 / bs^2) * 10) with 10 folded into 9."""
    rowptr, colind = _i32(rowptr), _i32(colind)
    n = rowptr.size - 1
    nb = (n + bs - 1) // bs
    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(rowptr))
    key = (rows // bs) * nb + colind.astype(np.int64) // bs
    counts = np.unique(key, return_counts=True)[1]
    occupy = (counts * 1.0 / (bs * bs)).astype(np.float32)
    ith = np.minimum((occupy * np.float32(10)).astype(np.int64), 9)
    return np.bincount(ith, minlength=10).astype(np.int64)


def ref_block_density_dist(rowptr, colind, bs):
    """The reference's own calculate_block_density_dist (oracle/_ref/libref_host.so)."""
    L = ref_host_lib()
    rowptr, colind = _i32(rowptr), _i32(colind)
    out = np.zeros(10, np.int64)
    L.ref_block_density_dist.argtypes = [C.c_int, _i32p, _i32p, C.c_int, _i64p]
    L.ref_block_density_dist(rowptr.size - 1, rowptr, colind, bs, out)
    return out


def ref_permutate(rowptr, colind, old2new):
    L = ref_host_lib()
    rowptr, colind, old2new = _i32(rowptr), _i32(colind), _i32(old2new)
    n = rowptr.size - 1
    nrp = np.zeros(n + 1, np.int32)
    nci = np.zeros(max(colind.size, 1), np.int32)
    L.ref_permutate.argtypes = [C.c_int, _i32p, _i32p, _i32p, _i32p, _i32p]
    L.ref_permutate(n, rowptr, colind, old2new, nrp, nci)
    return nrp, nci[:colind.size]


def ref_rcmk_graph(rowptr, colind):
    L = ref_host_lib()
    rowptr, colind = _i32(rowptr), _i32(colind)
    n = rowptr.size - 1
    nrp = np.zeros(n + 1, np.int32)
    nci = np.zeros(max(colind.size, 1), np.int32)
    L.ref_rcmk_graph.argtypes = [C.c_int, _i32p, _i32p, _i32p, _i32p]
    L.ref_rcmk_graph(n, rowptr, colind, nrp, nci)
    return nrp, nci[:colind.size]


def _ref_graph_call(name, rowptr, colind):
    L = ref_host_lib()
    rowptr, colind = _i32(rowptr), _i32(colind)
    n = rowptr.size - 1
    nrp = np.zeros(n + 1, np.int32)
    nci = np.zeros(max(colind.size, 1), np.int32)
    f = getattr(L, name)
    f.argtypes = [C.c_int, _i32p, _i32p, _i32p, _i32p]
    f(n, rowptr, colind, nrp, nci)
    return nrp, nci[:colind.size]


def ref_bfs_graph(rowptr, colind):
    """The reference's own BFSTraversal (reorder_strategy.cc:68-98): returns the REORDERED graph."""
    return _ref_graph_call("ref_bfs_graph", rowptr, colind)


def ref_max_degree_sort_graph(rowptr, colind):
    """The reference's own maxDegreeSort (reorder_strategy.cc:41-55): returns the REORDERED graph."""
    return _ref_graph_call("ref_max_degree_sort_graph", rowptr, colind)


_ref_divide = None


def ref_divide_lib():
    global _ref_divide
    p = os.path.join(HERE, "_ref", "libref_divide.so")
    if _ref_divide is None and os.path.exists(p):
        try:
            _ref_divide = C.CDLL(p)
        except OSError:
            return None
    return _ref_divide


def ref_divide_matrix(rowptr, colind, bs, density):
    """The reference's own divide_matrix (divide.cu:52-127): unit weights."""
    rowptr, colind = _i32(rowptr), _i32(colind)
    n = rowptr.size - 1
    nb = (n + bs - 1) // bs
    f = ref_divide_lib().ref_divide_matrix
    f.restype = C.c_longlong
    f.argtypes = [C.c_int, _i32p, _i32p, C.c_int, C.c_float] + [C.c_void_p] * 5 + [C.POINTER(C.c_longlong)]
    cnnz = C.c_longlong(0)
    nnzb = int(f(n, rowptr, colind, bs, density, None, None, None, None, None, C.byref(cnnz)))
    crp = np.zeros(n + 1, np.int32)
    cci = np.zeros(max(cnnz.value, 1), np.int32)
    brp = np.zeros(nb + 1, np.int32)
    bci = np.zeros(max(nnzb, 1), np.int32)
    bv = np.zeros(max(nnzb, 1) * bs * bs, np.float32)
    f(n, rowptr, colind, bs, density, crp.ctypes.data, cci.ctypes.data, brp.ctypes.data, bci.ctypes.data,
      bv.ctypes.data, C.byref(cnnz))
    return crp, cci[:cnnz.value], brp, bci[:nnzb], bv[:nnzb * bs * bs].reshape(nnzb, bs, bs)


# ---- graph files (load_data.cc:125-184), numpy restatement + the reference's own loaders ------------------
def dump_csr_text(prefix, rowptr, colind):
    """dumpCSRToFile (load_data.cc:125-141): "<count>\\n", every value followed by a space, "\\n"."""
    for name, a in (("_indptr.txt", rowptr), ("_indices.txt", colind)):
        with open(prefix + name, "w") as f:
            f.write(f"{len(a)}\n" + "".join(f"{int(x)} " for x in a) + "\n")


def load_csr_text(prefix):
    """loadCSRFromFile (load_data.cc:143-165): the first <count> integers after the count."""
    out = []
    for name in ("_indptr.txt", "_indices.txt"):
        tok = open(prefix + name).read().split()
        cnt = int(tok[0])
        out.append(np.array(tok[1:1 + cnt], dtype=np.int64).astype(np.int32))
    return out[0], out[1]


def load_edge_list_csr(path):
    """loadGraphFromFile (load_data.cc:167-184: push_back per pair, std::sort per row, duplicates kept,
    nnz from the header) + convertGraphToCSR (utility.cc:31-45)."""
    tok = open(path).read().split()
    n, nnz = int(tok[0]), int(tok[1])
    adj = [[] for _ in range(n)]
    for e in range(nnz):
        adj[int(tok[2 + 2 * e])].append(int(tok[3 + 2 * e]))
    rowptr = np.zeros(n + 1, np.int32)
    rowptr[1:] = np.cumsum([len(a) for a in adj])
    colind = np.array([c for a in adj for c in sorted(a)], np.int32)
    return rowptr, colind


def ref_load_csr(prefix, cap_rows, cap_nnz):
    L = ref_host_lib()
    L.ref_load_csr.argtypes = [C.c_char_p, _i32p, C.c_int, _i32p, C.c_int, C.POINTER(C.c_int)]
    rp, ci, nnz = np.zeros(cap_rows, np.int32), np.zeros(max(cap_nnz, 1), np.int32), C.c_int(0)
    n = L.ref_load_csr(prefix.encode(), rp, cap_rows, ci, cap_nnz, C.byref(nnz))
    return rp[:n + 1], ci[:nnz.value]


def ref_load_graph_csr(path, cap_rows, cap_nnz):
    L = ref_host_lib()
    L.ref_load_graph_csr.argtypes = [C.c_char_p, _i32p, C.c_int, _i32p, C.c_int, C.POINTER(C.c_int)]
    rp, ci, nnz = np.zeros(cap_rows, np.int32), np.zeros(max(cap_nnz, 1), np.int32), C.c_int(0)
    n = L.ref_load_graph_csr(path.encode(), rp, cap_rows, ci, cap_nnz, C.byref(nnz))
    return rp[:n + 1], ci[:nnz.value]


def ref_dump_csr(prefix, rowptr, colind):
    L = ref_host_lib()
    L.ref_dump_csr.argtypes = [C.c_char_p, C.c_int, C.c_int, _i32p, _i32p]
    L.ref_dump_csr(prefix.encode(), len(rowptr) - 1, len(colind), np.ascontiguousarray(rowptr, np.int32),
                   np.ascontiguousarray(colind if len(colind) else np.zeros(1), np.int32))

"""Host-side mirror of the reference's operator interface for the hot path,
on torch CUDA tensors (device memory + streams only; all arithmetic happens in
libspmmdb.so).  Names and argument meaning follow the reference:
gespmm_csrmm (gespmm_csrmm.h:422-426), rocsparse_bsrmm_template
(rocsparse_bsrmm.h:102-108), cusparseXcsr2bsrNnz / cusparseScsr2bsr
(run_bsrmm.cu:121-142), cusparseSbsr2csr (bsr2csr.cu:186-188), the per-block
GEMM loop of block_cublas.cu:123-136."""
import ctypes as C

import torch

from ._lib import lib, check

ROW_MAJOR, COL_MAJOR = 0, 1
DIR_ROW, DIR_COLUMN = 0, 1
MATH_FP32, MATH_TF32, MATH_BF16, MATH_BF16_PACKED = 0, 1, 2, 3


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else C.c_void_p(0)


def _dev(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise ValueError("spmmdb ops take CUDA tensors (there is no CPU path)")
        if t is not None and not t.is_contiguous():
            raise ValueError("spmmdb ops take contiguous tensors")


def csrmm_segment(nnz, m=None):
    """Segment length of a call: lower bound by nnz alone, or (with m) the length actually used."""
    if m is None:
        return lib().spmmdb_csrmm_segment(int(nnz))
    return lib().spmmdb_csrmm_segment_for(int(m), int(nnz))


def set_csrmm_segment(seg):
    """Process-wide segment override (0 = automatic); row-sharded runs set the whole matrix's length."""
    check(lib().spmmdb_set_csrmm_segment(int(seg)), "spmmdb_set_csrmm_segment")


def set_tuning(name, value=None):
    """Override a tuning knob (None clears it): spmmdb_set_tuning."""
    v = -2147483648 if value is None else int(value)
    check(lib().spmmdb_set_tuning(name.encode(), v), "spmmdb_set_tuning")


def get_tuning(name, default=0):
    return lib().spmmdb_get_tuning(name.encode(), int(default))


def csrmm(rowptr, colind, val, B, C_out, m, k, K, ldb=None, b_layout=ROW_MAJOR, ldc=None, c_layout=ROW_MAJOR,
          alpha=1.0, beta=0.0, nnz=None):
    """C = alpha*A*B + beta*C with explicit layouts (spmmdb_csrmm_f32)."""
    _dev(rowptr, colind, val, B, C_out)
    if ldb is None:
        ldb = K if b_layout == ROW_MAJOR else k
    if ldc is None:
        ldc = K if c_layout == ROW_MAJOR else m
    nnz = colind.numel() if nnz is None else nnz
    rc = lib().spmmdb_csrmm_f32(m, k, K, nnz, _p(rowptr), _p(colind), _p(val), _p(B), ldb, b_layout, _p(C_out), ldc,
                                c_layout, alpha, beta, _stream())
    check(rc, "spmmdb_csrmm_f32")
    return C_out


def csr_hot_columns(k, colind, max_hot_columns):
    """Bitmap (int32 words) of the most frequent columns that fit max_hot_columns, and how many were marked."""
    _dev(colind)
    bitmap = torch.zeros((k + 31) // 32, dtype=torch.int32, device=colind.device)
    n_hot = C.c_int(0)
    rc = lib().spmmdb_csr_hot_columns(k, colind.numel(), _p(colind), int(max_hot_columns), _p(bitmap), C.byref(n_hot),
                                      _stream())
    check(rc, "spmmdb_csr_hot_columns")
    return bitmap, n_hot.value


def dense_pack_bf16(B, K=None):
    """fp32 [rows, K] -> bf16 [rows, K] (spmmdb_dense_pack_bf16, round to nearest even)."""
    _dev(B)
    rows, ld = B.shape
    K = ld if K is None else K
    out = torch.empty(rows, K, dtype=torch.bfloat16, device=B.device)
    check(lib().spmmdb_dense_pack_bf16(rows, K, _p(B), ld, _p(out), K, _stream()), "spmmdb_dense_pack_bf16")
    return out


def csrmm_ex(rowptr, colind, val, B, C_out, m, k, K, ldb=None, ldc=None, alpha=1.0, beta=0.0, hot=None, nnz=None):
    """spmmdb_csrmm_ex_f32: row-major, B fp32 or bf16 (by tensor dtype), optional hot-column bitmap."""
    _dev(rowptr, colind, val, B, C_out, hot)
    dtype = 1 if B.dtype == torch.bfloat16 else 0
    nnz = colind.numel() if nnz is None else nnz
    rc = lib().spmmdb_csrmm_ex_f32(m, k, K, nnz, _p(rowptr), _p(colind), _p(val), _p(B), dtype, K if ldb is None else ldb,
                                   _p(C_out), K if ldc is None else ldc, alpha, beta, _p(hot), _stream())
    check(rc, "spmmdb_csrmm_ex_f32")
    return C_out


def csrmm_noalloc(rowptr, colind, val, B, C_out, m, k, K, ldb=None, ldc=None, alpha=1.0, beta=0.0):
    """spmmdb_csrmm_noalloc_f32: row-major CSR x dense whose scratch is a fixed device array inside the library
    (no cudaMalloc in the call; K <= 512; calls that overlap on the GPU race on the array)."""
    _dev(rowptr, colind, val, B, C_out)
    rc = lib().spmmdb_csrmm_noalloc_f32(m, k, K, _p(rowptr), _p(colind), _p(val), _p(B), K if ldb is None else ldb,
                                        _p(C_out), K if ldc is None else ldc, alpha, beta, _stream())
    check(rc, "spmmdb_csrmm_noalloc_f32")
    return C_out


def gespmm_csrmm(A_nrows, B_ncols, A_rowPtr, A_colInd, A_val, B, C_out):
    """gespmm_csrmm<float>(A_nrows, B_ncols, rowPtr, colInd, val, B, C): row-major B/C,
    alpha=1, beta=0, nnz unknown to the callee (gespmm_csrmm.h:422-426)."""
    _dev(A_rowPtr, A_colInd, A_val, B, C_out)
    k = B.numel() // max(B_ncols, 1)
    rc = lib().spmmdb_csrmm_f32(A_nrows, k, B_ncols, -1, _p(A_rowPtr), _p(A_colInd), _p(A_val), _p(B), B_ncols,
                                ROW_MAJOR, _p(C_out), B_ncols, ROW_MAJOR, 1.0, 0.0, _stream())
    check(rc, "gespmm_csrmm")
    return C_out


def bsrmm(direction, trans_b, mb, n, kb, nnzb, alpha, bsr_val, bsr_rowptr, bsr_colind, block_dim, B, ldb, beta,
          C_out, ldc, c_layout=COL_MAJOR, math=MATH_FP32):
    _dev(bsr_val, bsr_rowptr, bsr_colind, B, C_out)
    rc = lib().spmmdb_bsrmm_f32(direction, int(trans_b), mb, n, kb, nnzb, block_dim, alpha, _p(bsr_val),
                                _p(bsr_rowptr), _p(bsr_colind), _p(B), ldb, beta, _p(C_out), ldc, c_layout, math,
                                _stream())
    check(rc, "spmmdb_bsrmm_f32")
    return C_out


def rocsparse_bsrmm(direction, trans_A, trans_B, mb, n, kb, nnzb, alpha, bsr_val, bsr_row_ptr, bsr_col_ind,
                    block_dim, B, ldb, beta, C_out, ldc, math=MATH_FP32):
    """rocsparse_bsrmm_template<float> argument order (rocsparse_bsrmm.h:102-108); returns the
    status code instead of raising, like the reference (0 = CUSPARSE_STATUS_SUCCESS)."""
    if trans_A != 0:
        return -2  # CUSPARSE_STATUS_MATRIX_TYPE_NOT_SUPPORTED (rocsparse_bsrmm.h:139-141)
    return lib().spmmdb_bsrmm_f32(direction, int(trans_B), mb, n, kb, nnzb, block_dim, alpha, _p(bsr_val),
                                  _p(bsr_row_ptr), _p(bsr_col_ind), _p(B), ldb, beta, _p(C_out), ldc, COL_MAJOR,
                                  math, _stream())


def bsr_transpose_blocks(bsr_val, block_dim):
    """Column-major blocks -> row-major blocks (or back), once per matrix: spmmdb_bsr_transpose_blocks_f32."""
    _dev(bsr_val)
    out = torch.empty_like(bsr_val)
    nnzb = bsr_val.numel() // (block_dim * block_dim)
    check(lib().spmmdb_bsr_transpose_blocks_f32(nnzb, block_dim, _p(bsr_val), _p(out), _stream()),
          "spmmdb_bsr_transpose_blocks_f32")
    return out


def block_gemm(mb, dim, nnzb, block_dim, bsr_rowptr, bsr_colind, bsr_val, Y, Z, math=MATH_FP32):
    _dev(bsr_rowptr, bsr_colind, bsr_val, Y, Z)
    rc = lib().spmmdb_block_gemm_f32(mb, dim, nnzb, block_dim, _p(bsr_rowptr), _p(bsr_colind), _p(bsr_val), _p(Y),
                                     _p(Z), math, _stream())
    check(rc, "spmmdb_block_gemm_f32")
    return Z


def csr2bsr_nnz(m, n, rowptr, colind, block_dim, bsr_rowptr):
    _dev(rowptr, colind, bsr_rowptr)
    nnzb = C.c_int(0)
    rc = lib().spmmdb_csr2bsr_nnz(m, n, _p(rowptr), _p(colind), block_dim, _p(bsr_rowptr), C.byref(nnzb), _stream())
    check(rc, "spmmdb_csr2bsr_nnz")
    return nnzb.value


def csr2bsr(m, n, rowptr, colind, val, block_dim, direction=DIR_ROW):
    """Both phases, allocating the outputs between them like run_bsrmm.cu:121-142."""
    mb = (m + block_dim - 1) // block_dim
    bsr_rowptr = torch.empty(mb + 1, dtype=torch.int32, device=rowptr.device)
    nnzb = csr2bsr_nnz(m, n, rowptr, colind, block_dim, bsr_rowptr)
    bsr_colind = torch.empty(nnzb, dtype=torch.int32, device=rowptr.device)
    bsr_val = torch.empty(nnzb * block_dim * block_dim, dtype=torch.float32, device=rowptr.device)
    rc = lib().spmmdb_csr2bsr_f32(direction, m, n, _p(rowptr), _p(colind), _p(val), block_dim, _p(bsr_rowptr),
                                  _p(bsr_colind), _p(bsr_val), _stream())
    check(rc, "spmmdb_csr2bsr_f32")
    return bsr_rowptr, bsr_colind, bsr_val


def bsr2csr(mb, nb, bsr_rowptr, bsr_colind, bsr_val, block_dim, direction=DIR_ROW):
    _dev(bsr_rowptr, bsr_colind, bsr_val)
    nnz = bsr_colind.numel() * block_dim * block_dim
    dev = bsr_rowptr.device
    csr_rowptr = torch.empty(mb * block_dim + 1, dtype=torch.int32, device=dev)
    csr_colind = torch.empty(nnz, dtype=torch.int32, device=dev)
    csr_val = torch.empty(nnz, dtype=torch.float32, device=dev)
    rc = lib().spmmdb_bsr2csr_f32(direction, mb, nb, _p(bsr_rowptr), _p(bsr_colind), _p(bsr_val), block_dim,
                                  _p(csr_rowptr), _p(csr_colind), _p(csr_val), _stream())
    check(rc, "spmmdb_bsr2csr_f32")
    return csr_rowptr, csr_colind, csr_val


def divide_threshold(block_dim, density):
    return lib().spmmdb_divide_threshold(block_dim, density)


def divide(m, n, rowptr, colind, val, block_dim, density, direction=DIR_ROW):
    """divide_matrix (divide.cu:52-127) on the device: blocks with fill >= density -> BSR, the rest ->
    remainder CSR.  Returns (csr_rowptr, csr_colind, csr_val, bsr_rowptr, bsr_colind, bsr_val)."""
    _dev(rowptr, colind, val)
    dev = rowptr.device
    mb = (m + block_dim - 1) // block_dim
    bsr_rowptr = torch.empty(mb + 1, dtype=torch.int32, device=dev)
    csr_rowptr = torch.empty(m + 1, dtype=torch.int32, device=dev)
    nnzb, cnnz = C.c_int(0), C.c_int(0)
    rc = lib().spmmdb_divide_nnz(m, n, _p(rowptr), _p(colind), block_dim, density, _p(bsr_rowptr), _p(csr_rowptr),
                                 C.byref(nnzb), C.byref(cnnz), _stream())
    check(rc, "spmmdb_divide_nnz")
    bsr_colind = torch.empty(nnzb.value, dtype=torch.int32, device=dev)
    bsr_val = torch.empty(nnzb.value * block_dim * block_dim, dtype=torch.float32, device=dev)
    csr_colind = torch.empty(cnnz.value, dtype=torch.int32, device=dev)
    csr_val = torch.empty(cnnz.value, dtype=torch.float32, device=dev)
    rc = lib().spmmdb_divide_f32(direction, m, n, _p(rowptr), _p(colind), _p(val), block_dim, density,
                                 _p(bsr_rowptr), _p(bsr_colind), _p(bsr_val), _p(csr_rowptr), _p(csr_colind),
                                 _p(csr_val), _stream())
    check(rc, "spmmdb_divide_f32")
    return csr_rowptr, csr_colind, csr_val, bsr_rowptr, bsr_colind, bsr_val


class Hybrid:
    """spmmdb_hybrid: A split once into dense blocks (tensor-core BSR) + remainder (CSR); spmm() multiplies both
    parts into one C.  The recommended BSR entry for (re-ordered) graph adjacency matrices."""

    def __init__(self, m, n, rowptr, colind, val, block_dim, density):
        _dev(rowptr, colind, val)
        self._h = C.c_void_p(0)
        rc = lib().spmmdb_hybrid_create(m, n, _p(rowptr), _p(colind), _p(val), block_dim, density, C.byref(self._h),
                                        _stream())
        check(rc, "spmmdb_hybrid_create")
        a, b = C.c_int64(0), C.c_int64(0)
        lib().spmmdb_hybrid_info(self._h, C.byref(a), C.byref(b))
        self.m, self.n, self.block_dim, self.nnzb, self.csr_nnz = m, n, block_dim, a.value, b.value

    def spmm(self, B, C_out, K, ldb=None, ldc=None, alpha=1.0, beta=0.0, math=MATH_TF32):
        _dev(B, C_out)
        rc = lib().spmmdb_hybrid_spmm_f32(self._h, K, _p(B), K if ldb is None else ldb, _p(C_out),
                                          K if ldc is None else ldc, alpha, beta, math, _stream())
        check(rc, "spmmdb_hybrid_spmm_f32")
        return C_out

    def close(self):
        if self._h:
            lib().spmmdb_hybrid_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def permute_csr(rowptr, colind, val, old2new):
    """A' = P A P^T (permutate, reorder_strategy.cc:26-39) on the device.  Returns (rowptr, colind, val)."""
    _dev(rowptr, colind, val, old2new)
    n = rowptr.numel() - 1
    nrp = torch.empty(n + 1, dtype=torch.int32, device=rowptr.device)
    nci = torch.empty(colind.numel(), dtype=torch.int32, device=rowptr.device)
    nv = torch.empty(colind.numel(), dtype=torch.float32, device=rowptr.device) if val is not None else None
    rc = lib().spmmdb_permute_csr_f32(n, _p(rowptr), _p(colind), _p(val), _p(old2new), _p(nrp), _p(nci), _p(nv),
                                      _stream())
    check(rc, "spmmdb_permute_csr_f32")
    return nrp, nci, nv


def order_by_degree(rowptr):
    """old2new of maxDegreeSort (reorder_strategy.cc:41-55): descending degree, ties by ascending id."""
    _dev(rowptr)
    n = rowptr.numel() - 1
    old2new = torch.empty(n, dtype=torch.int32, device=rowptr.device)
    check(lib().spmmdb_order_by_degree(n, _p(rowptr), _p(old2new), _stream()), "spmmdb_order_by_degree")
    return old2new


def order_bfs(rowptr, colind):
    """old2new of BFSTraversal (reorder_strategy.cc:68-98); colind in adjacency order (need not be sorted)."""
    _dev(rowptr, colind)
    n = rowptr.numel() - 1
    old2new = torch.empty(n, dtype=torch.int32, device=rowptr.device)
    check(lib().spmmdb_order_bfs(n, _p(rowptr), _p(colind), _p(old2new), _stream()), "spmmdb_order_bfs")
    return old2new


def order_rcm(rowptr, colind):
    """old2new of reverseCuthillMcKee (reorder_strategy.cc:57-66): lists by descending degree (ties: ascending id), BFS."""
    _dev(rowptr, colind)
    n = rowptr.numel() - 1
    old2new = torch.empty(n, dtype=torch.int32, device=rowptr.device)
    check(lib().spmmdb_order_rcm(n, _p(rowptr), _p(colind), _p(old2new), _stream()), "spmmdb_order_rcm")
    return old2new


def permute_rows(X, old2new):
    """out[old2new[r]] = X[r] for a row-major dense matrix."""
    _dev(X, old2new)
    out = torch.empty_like(X)
    rows, K = X.shape
    check(lib().spmmdb_permute_rows_f32(rows, K, _p(X), K, _p(old2new), _p(out), K, _stream()),
          "spmmdb_permute_rows_f32")
    return out


def heatmap(rowptr, colind, block_dim=256):
    """getHeatmap (utility.cc:71-88): non-zeros per block_dim x block_dim block, nb x nb int32."""
    _dev(rowptr, colind)
    n = rowptr.numel() - 1
    nb = (n + block_dim - 1) // block_dim
    heat = torch.empty(nb, nb, dtype=torch.int32, device=rowptr.device)
    check(lib().spmmdb_heatmap(n, _p(rowptr), _p(colind), block_dim, _p(heat), _stream()), "spmmdb_heatmap")
    return heat


def block_fill_histogram(rowptr, colind, block_dim, n=None):
    """calculate_block_density_dist (block_density_dist.cpp:47-86): blocks per fill decile, as a list of 10 ints."""
    _dev(rowptr, colind)
    m = rowptr.numel() - 1
    dist = (C.c_int64 * 10)()
    rc = lib().spmmdb_block_fill_histogram(m, m if n is None else n, _p(rowptr), _p(colind), block_dim, dist, _stream())
    check(rc, "spmmdb_block_fill_histogram")
    return list(dist)


def block_metrics(rowptr, colind, block_dim):
    """analyzeBlockSparseMetrics (reorder_graph.cc:12-24): (nnzb, density, utilization, average)."""
    n, nnz = rowptr.numel() - 1, colind.numel()
    nb = (n + block_dim - 1) // block_dim
    tmp = torch.empty(nb + 1, dtype=torch.int32, device=rowptr.device)
    nnzb = csr2bsr_nnz(n, n, rowptr, colind, block_dim, tmp)
    return nnzb, nnzb / (nb * nb), nnz / (nnzb * block_dim * block_dim), nnz / nnzb


def bsr_pack_bf16(bsr_val, block_dim):
    """bf16 copy of the blocks for bsrmm(..., math=MATH_BF16_PACKED): made once per matrix."""
    _dev(bsr_val)
    out = torch.empty(bsr_val.numel(), dtype=torch.bfloat16, device=bsr_val.device)
    nnzb = bsr_val.numel() // (block_dim * block_dim)
    check(lib().spmmdb_bsr_pack_bf16(nnzb, block_dim, _p(bsr_val), _p(out), _stream()), "spmmdb_bsr_pack_bf16")
    return out

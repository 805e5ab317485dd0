"""ctypes binding of libspmmdb.so (include/spmmdb.h).  Fails loudly when the
library is missing -- the product never falls back to another implementation."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_i64 = C.c_int64
vp = C.c_void_p


class SpmmdbError(RuntimeError):
    pass


STATUS = {0: "OK", -1: "INVALID_VALUE", -2: "NOT_SUPPORTED", -3: "CUDA", -4: "ALLOC", -5: "ARCH", -6: "IO"}

# name -> argtypes; every symbol include/spmmdb.h declares
SIGNATURES = {
    "spmmdb_csrmm_f32": [C.c_int, C.c_int, C.c_int, c_i64, vp, vp, vp, vp, c_i64, C.c_int, vp, c_i64, C.c_int,
                         C.c_float, C.c_float, vp],
    "spmmdb_csrmm_ex_f32": [C.c_int, C.c_int, C.c_int, c_i64, vp, vp, vp, vp, C.c_int, c_i64, vp, c_i64, C.c_float,
                            C.c_float, vp, vp],
    "spmmdb_csrmm_noalloc_f32": [C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, c_i64, vp, c_i64, C.c_float, C.c_float, vp],
    "spmmdb_order_by_degree": [C.c_int, vp, vp, vp],
    "spmmdb_order_bfs": [C.c_int, vp, vp, vp, vp],
    "spmmdb_order_rcm": [C.c_int, vp, vp, vp, vp],
    "spmmdb_csr_hot_columns": [C.c_int, c_i64, vp, c_i64, vp, C.POINTER(C.c_int), vp],
    "spmmdb_dense_pack_bf16": [c_i64, C.c_int, vp, c_i64, vp, c_i64, vp],
    "spmmdb_csrmm_segment": [c_i64],
    "spmmdb_csrmm_segment_for": [C.c_int, c_i64],
    "spmmdb_set_csrmm_segment": [C.c_int],
    "spmmdb_get_csrmm_segment": [],
    "spmmdb_bsrmm_f32": [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, c_i64, C.c_int, C.c_float, vp, vp, vp, vp,
                         c_i64, C.c_float, vp, c_i64, C.c_int, C.c_int, vp],
    "spmmdb_bsr_pack_bf16": [c_i64, C.c_int, vp, vp, vp],
    "spmmdb_bsr_transpose_blocks_f32": [c_i64, C.c_int, vp, vp, vp],
    "spmmdb_block_gemm_f32": [C.c_int, C.c_int, c_i64, C.c_int, vp, vp, vp, vp, vp, C.c_int, vp],
    "spmmdb_csr2bsr_nnz": [C.c_int, C.c_int, vp, vp, C.c_int, vp, C.POINTER(C.c_int), vp],
    "spmmdb_csr2bsr_f32": [C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int, vp, vp, vp, vp],
    "spmmdb_bsr2csr_f32": [C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int, vp, vp, vp, vp],
    "spmmdb_divide_threshold": [C.c_int, C.c_float],
    "spmmdb_divide_nnz": [C.c_int, C.c_int, vp, vp, C.c_int, C.c_float, vp, vp, C.POINTER(C.c_int),
                          C.POINTER(C.c_int), vp],
    "spmmdb_divide_f32": [C.c_int, C.c_int, C.c_int, vp, vp, vp, C.c_int, C.c_float, vp, vp, vp, vp, vp, vp, vp],
    "spmmdb_hybrid_create": [C.c_int, C.c_int, vp, vp, vp, C.c_int, C.c_float, C.POINTER(vp), vp],
    "spmmdb_hybrid_info": [vp, C.POINTER(c_i64), C.POINTER(c_i64)],
    "spmmdb_hybrid_spmm_f32": [vp, C.c_int, vp, c_i64, vp, c_i64, C.c_float, C.c_float, C.c_int, vp],
    "spmmdb_hybrid_destroy": [vp],
    "spmmdb_permute_csr_f32": [C.c_int, vp, vp, vp, vp, vp, vp, vp, vp],
    "spmmdb_permute_rows_f32": [C.c_int, C.c_int, vp, c_i64, vp, vp, c_i64, vp],
    "spmmdb_block_fill_histogram": [C.c_int, C.c_int, vp, vp, C.c_int, C.POINTER(c_i64), vp],
    "spmmdb_heatmap": [C.c_int, vp, vp, C.c_int, vp, vp],
    "spmmdb_csrmm_host_f32": [C.c_int, C.c_int, C.c_int, c_i64, vp, vp, vp, vp, vp, C.c_float, C.c_float, vp],
    "spmmdb_set_bsrmm_math": [C.c_int],
    "spmmdb_get_bsrmm_math": [],
    "spmmdb_shard_plan": [vp, C.c_int, C.c_int, C.c_int, vp],
    "spmmdb_mg_unique_id": [vp],
    "spmmdb_mg_create": [C.c_int, C.c_int, vp, C.POINTER(vp)],
    "spmmdb_mg_create_from_comm": [vp, C.c_int, C.c_int, C.POINTER(vp)],
    "spmmdb_mg_destroy": [vp],
    "spmmdb_mg_allgather_rows_f32": [vp, vp, c_i64, C.c_int, vp, vp],
    "spmmdb_sharded_csrmm_f32": [vp, C.c_int, C.c_int, c_i64, vp, vp, vp, vp, c_i64, vp, vp, c_i64, C.c_float,
                                 C.c_float, C.c_int, vp],
    "spmmdb_graph_read": [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(c_i64), C.POINTER(vp), C.POINTER(vp)],
    "spmmdb_graph_write": [C.c_char_p, C.c_int, C.c_int, c_i64, vp, vp],
    "spmmdb_graph_free": [vp],
    "spmmdb_set_tuning": [C.c_char_p, C.c_int],
    "spmmdb_get_tuning": [C.c_char_p, C.c_int],
    "spmmdb_version": [],
    "spmmdb_last_error": [],
    "spmmdb_device_sm": [],
}


def lib_path():
    return os.path.join(HERE, "libspmmdb.so")


def lib():
    global _LIB
    if _LIB is None:
        p = lib_path()
        if not os.path.exists(p):
            raise SpmmdbError(
                f"{p} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C spmm_denseblock_b200/csrc). There is no fallback path.")
        L = C.CDLL(p)
        for name, argtypes in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the ABI lost a symbol
            fn.argtypes = argtypes
            fn.restype = C.c_int
        L.spmmdb_version.restype = C.c_char_p
        L.spmmdb_last_error.restype = C.c_char_p
        L.spmmdb_graph_free.restype = None
        _LIB = L
    return _LIB


def check(rc, what):
    if rc != 0:
        msg = lib().spmmdb_last_error().decode() if rc == -3 else ""
        raise SpmmdbError(f"{what} failed: {STATUS.get(rc, rc)} {msg}")

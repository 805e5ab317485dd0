"""spmm_denseblock_b200 -- B200-native SpMM engine (CSR / BSR x dense, csr2bsr,
bsr2csr) behind the reference's operator interface.

The product is libspmmdb.so (hand-written sm_100a CUDA behind the C ABI of
include/spmmdb.h).  This package is the thin host-side mirror used by the tests
and the benchmark: torch supplies device memory, streams and torch.distributed,
nothing else.  There is no CPU fallback: importing works anywhere, calling any
op without the built library or without a CUDA device raises.
"""
from ._lib import lib, lib_path, SpmmdbError  # noqa: F401
from .ops import (  # noqa: F401
    ROW_MAJOR, COL_MAJOR, DIR_ROW, DIR_COLUMN, MATH_FP32, MATH_TF32, MATH_BF16, MATH_BF16_PACKED, bsr_pack_bf16,
    gespmm_csrmm, csrmm, bsrmm, rocsparse_bsrmm, block_gemm, bsr_transpose_blocks, csr2bsr_nnz, csr2bsr, bsr2csr,
    csrmm_segment, set_csrmm_segment, set_tuning, get_tuning, csrmm_ex, csrmm_noalloc, csr_hot_columns, dense_pack_bf16, divide, divide_threshold, permute_csr, permute_rows, order_by_degree, order_bfs, order_rcm, heatmap, block_metrics, block_fill_histogram, Hybrid,
)
from .graphio import (  # noqa: F401
    GRAPH_TEXT, GRAPH_BINARY, GRAPH_EDGELIST, GRAPH_AUTO, read_graph, write_graph,
)

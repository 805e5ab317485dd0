"""Graph files of the reference's pipeline (load_data.cc:125-184) plus a binary cache format, through
libspmmdb.so's host-side readers (csrc/graphio.cu; no CUDA call, works without a GPU).  Arrays come back as
pinned-or-pageable CPU int32 tensors ready for .cuda()."""
import ctypes as C

import numpy as np
import torch

from ._lib import lib, check

GRAPH_TEXT, GRAPH_BINARY, GRAPH_EDGELIST, GRAPH_AUTO = 0, 1, 2, 3


def read_graph(path, format=GRAPH_AUTO, pin=False):
    """(rowptr, colind) of `<path>_indptr.txt/_indices.txt` (loadCSRFromFile), an edge-list file
    (loadGraphFromFile + convertGraphToCSR) or `<path>.csrbin`; AUTO prefers the binary file."""
    n, nnz = C.c_int(0), C.c_int64(0)
    rp, ci = C.c_void_p(0), C.c_void_p(0)
    L = lib()
    check(L.spmmdb_graph_read(str(path).encode(), format, C.byref(n), C.byref(nnz), C.byref(rp), C.byref(ci)),
          "spmmdb_graph_read")
    try:
        rowptr = torch.from_numpy(np.ctypeslib.as_array(C.cast(rp, C.POINTER(C.c_int)), (n.value + 1,)).copy())
        colind = (torch.from_numpy(np.ctypeslib.as_array(C.cast(ci, C.POINTER(C.c_int)), (nnz.value,)).copy())
                  if nnz.value else torch.empty(0, dtype=torch.int32))
    finally:
        L.spmmdb_graph_free(rp)
        L.spmmdb_graph_free(ci)
    if pin:
        rowptr, colind = rowptr.pin_memory(), colind.pin_memory()
    return rowptr, colind


def write_graph(prefix, rowptr, colind, format=GRAPH_BINARY):
    """dumpCSRToFile's text pair (byte-identical) or the binary cache file."""
    rp = rowptr.detach().cpu().to(torch.int32).contiguous()
    ci = colind.detach().cpu().to(torch.int32).contiguous()
    check(lib().spmmdb_graph_write(str(prefix).encode(), format, rp.numel() - 1, ci.numel(),
                                   C.c_void_p(rp.data_ptr()), C.c_void_p(ci.data_ptr() if ci.numel() else 0)),
          "spmmdb_graph_write")

"""nnz-balanced row sharding of a CSR matrix over the GPUs of one box (BASELINE.json
north_star; the reference itself is single-GPU).  Output rows are disjoint, so the SpMM
needs no reduction; B is replicated (all-gather of the row shards of B)."""
import torch


def nnz_balanced_ranges(rowptr, parts):
    """Contiguous row ranges [lo_g, hi_g) holding ~nnz/parts non-zeros each:
    boundary g = first row whose rowptr >= g*nnz/parts."""
    n = rowptr.numel() - 1
    nnz = int(rowptr[-1].item())
    targets = torch.tensor([(g * nnz) // parts for g in range(1, parts)], dtype=rowptr.dtype, device=rowptr.device)
    cuts = torch.searchsorted(rowptr, targets, right=False).clamp_(0, n).tolist() if parts > 1 else []
    bounds = [0] + cuts + [n]
    for i in range(1, len(bounds)):
        bounds[i] = max(bounds[i], bounds[i - 1])
    return [(bounds[g], bounds[g + 1]) for g in range(parts)]


def shard_csr(rowptr, colind, val, lo, hi):
    """Row slice [lo, hi) with rowptr rebased to 0; column ids stay global."""
    s, e = int(rowptr[lo].item()), int(rowptr[hi].item())
    rp = (rowptr[lo:hi + 1] - rowptr[lo]).contiguous()
    return rp, colind[s:e].contiguous(), (val[s:e].contiguous() if val is not None else None)


def shard_plan(rowptr, parts, align=1):
    """Same ranges through the C ABI (spmmdb_shard_plan: host function, no CUDA call)."""
    import ctypes as C

    import numpy as np

    from ._lib import lib, check
    rp = np.ascontiguousarray(rowptr.detach().cpu().numpy() if hasattr(rowptr, "detach") else rowptr, dtype=np.int32)
    bounds = np.zeros(parts + 1, np.int32)
    check(lib().spmmdb_shard_plan(rp.ctypes.data_as(C.c_void_p), rp.size - 1, parts, align,
                                  bounds.ctypes.data_as(C.c_void_p)), "spmmdb_shard_plan")
    return [(int(bounds[g]), int(bounds[g + 1])) for g in range(parts)]


def allgather_rows(B, ranges, group=None):
    """Replicate B when rank g owns rows ranges[g] of it (e.g. the C rows it produced as the previous
    layer's output): an all-gather-v as one broadcast per owner, in place into the full-size B.
    Row counts differ per rank (nnz-balanced ranges), which rules out a plain all_gather.
    Works on NCCL (device tensors, NVLink) and gloo (CPU tests)."""
    import torch.distributed as dist
    works = []
    for g, (lo, hi) in enumerate(ranges):
        if hi > lo:
            works.append(dist.broadcast(B[lo:hi], src=g, group=group, async_op=True))
    for w in works:
        w.wait()
    return B


class MultiGpu:
    """spmmdb_mg handle of this rank (one process per GPU): NCCL communicator + the library's communication
    stream, made collectively.  The 128-byte NCCL id travels over torch.distributed (any backend); everything
    after that is the C ABI (spmmdb_mg_*, spmmdb_sharded_csrmm_f32)."""

    def __init__(self, rank=None, world=None, group=None):
        import ctypes as C

        import torch.distributed as dist

        from ._lib import lib, check
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world is None else world
        L = lib()
        ident = [None]
        if self.rank == 0:
            buf = C.create_string_buffer(128)
            check(L.spmmdb_mg_unique_id(buf), "spmmdb_mg_unique_id")
            ident[0] = bytes(buf.raw)
        if self.world > 1:
            dist.broadcast_object_list(ident, src=0, group=group)
        self._h = C.c_void_p(0)
        check(L.spmmdb_mg_create(self.rank, self.world, C.c_char_p(ident[0]), C.byref(self._h)), "spmmdb_mg_create")
        self._bounds = None

    def _b(self, ranges):
        import numpy as np
        b = np.asarray([r[0] for r in ranges] + [ranges[-1][1]], np.int32)
        self._bounds = b  # keep alive across the asynchronous call
        return b.ctypes.data

    def allgather_rows(self, B, ranges):
        """In-place all-gather-v of the row ranges of a full-size row-major B (one NCCL send / recv group)."""
        import ctypes as C

        import torch

        from ._lib import lib, check
        check(lib().spmmdb_mg_allgather_rows_f32(self._h, C.c_void_p(B.data_ptr()), B.stride(0), B.shape[0],
                                                 C.c_void_p(self._b(ranges)),
                                                 C.c_void_p(torch.cuda.current_stream().cuda_stream)),
              "spmmdb_mg_allgather_rows_f32")
        return B

    def sharded_csrmm(self, rowptr, colind, val, B_local, ranges, C_out, k, K, alpha=1.0, beta=0.0, slabs=0):
        """This rank's rows of C = A * B, from its shard of A and ITS rows of B (spmmdb_sharded_csrmm_f32)."""
        import ctypes as C

        import torch

        from ._lib import lib, check
        p = lambda t: C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else C.c_void_p(0)  # noqa: E731
        check(lib().spmmdb_sharded_csrmm_f32(self._h, k, K, colind.numel(), p(rowptr), p(colind), p(val), p(B_local),
                                             B_local.stride(0) if B_local.numel() else K, C.c_void_p(self._b(ranges)),
                                             p(C_out), C_out.stride(0) if C_out.numel() else K, alpha, beta, slabs,
                                             C.c_void_p(torch.cuda.current_stream().cuda_stream)),
              "spmmdb_sharded_csrmm_f32")
        return C_out

    def close(self):
        from ._lib import lib
        if self._h:
            lib().spmmdb_mg_destroy(self._h)
            self._h = None

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

"""Multi-GPU row sharding: the plan, the shard slicing, the all-gather-v of B and
the "concatenated shards == single-device result" property.  The N > 1 path is exercised here on
the CPU with gloo, world size 2 (each rank's SpMM is done by the oracle -- there is no GPU in this
leg and the product has no CPU path); the GPU leg runs the real kernel shard by shard."""
import os
import socket

import numpy as np
import pytest
import torch

from helpers import random_csr


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_plan_matches_and_balances():
    from spmm_denseblock_b200 import shard
    rng = np.random.default_rng(3)
    rp, ci, _ = random_csr(rng, 5000, 4000, 9, hubs={17: 900, 4000: 1500}, empty_rows=range(100, 160))
    rpt = torch.from_numpy(rp)
    for parts in (1, 2, 3, 4, 8):
        a = shard.nnz_balanced_ranges(rpt, parts)
        b = shard.shard_plan(rpt, parts)
        assert a == b
        assert a[0][0] == 0 and a[-1][1] == 5000
        assert all(a[i][1] == a[i + 1][0] for i in range(parts - 1))
        nnz = [int(rp[hi] - rp[lo]) for lo, hi in a]
        assert sum(nnz) == rp[-1]
        # every shard within one (longest) row of the ideal share
        assert max(nnz) <= rp[-1] / parts + np.diff(rp).max()
    # block-aligned boundaries for a matrix that is cut before csr2bsr
    c = shard.shard_plan(rpt, 4, align=16)
    assert all(lo % 16 == 0 for lo, _ in c) and c[-1][1] == 5000
    # degenerate inputs
    z = torch.zeros(11, dtype=torch.int32)
    assert shard.shard_plan(z, 3) == [(0, 0), (0, 0), (0, 10)]
    from spmm_denseblock_b200._lib import lib
    assert lib().spmmdb_shard_plan(None, 3, 2, 1, None) == -1


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from oracle import oracle as O
    from spmm_denseblock_b200 import shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(11)  # same matrix on every rank
    n, K = 1500, 24
    rp, ci, v = random_csr(rng, n, n, 6, hubs={5: 700})
    Bfull = rng.uniform(-1, 1, (n, K)).astype(np.float32)
    ranges = shard.shard_plan(torch.from_numpy(rp), world)
    lo, hi = ranges[rank]
    # this rank owns only its rows of B; everything else is garbage until the all-gather
    B = torch.full((n, K), float("nan"))
    B[lo:hi] = torch.from_numpy(Bfull[lo:hi])
    shard.allgather_rows(B, ranges)
    assert torch.equal(B, torch.from_numpy(Bfull))
    rps, cis, vs = shard.shard_csr(torch.from_numpy(rp), torch.from_numpy(ci), torch.from_numpy(v), lo, hi)
    Cs = O.csrmm(rps.numpy(), cis.numpy(), vs.numpy(), B.numpy(), K)
    parts = [None] * world
    dist.all_gather_object(parts, (lo, hi, np.asarray(Cs)))
    if rank == 0:
        full = O.csrmm(rp, ci, v, Bfull, K)
        cat = np.concatenate([c for _, _, c in sorted(parts, key=lambda t: t[0])])
        out.put(bool(np.array_equal(cat, np.asarray(full))))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_spmm_gloo_world2(oracle):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    assert out.get(timeout=5) is True


@pytest.mark.gpu
@pytest.mark.parametrize("parts", [2, 3, 8])
def test_shards_concatenate_bit_exact_gpu(parts):
    """Rows longer than the segment length are split per row, not per shard: concatenated shard
    outputs are bit-identical to the single-GPU output (fp32 CSR path)."""
    import spmm_denseblock_b200 as pkg
    from spmm_denseblock_b200 import shard
    rng = np.random.default_rng(21)
    n, K = 6000, 128
    rp, ci, v = random_csr(rng, n, n, 12, hubs={7: 3000, 4100: 2500}, empty_rows=range(50, 70))
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    rpd, cid, vd = d(rp), d(ci), d(v)
    B = d(rng.uniform(-1, 1, (n, K)).astype(np.float32))
    nnz = ci.size
    full = torch.empty(n, K, device="cuda")
    pkg.csrmm(rpd, cid, vd, B, full, n, n, K, nnz=nnz)
    outs = []
    for lo, hi in shard.shard_plan(rpd, parts):
        rps, cis, vs = shard.shard_csr(rpd, cid, vd, lo, hi)
        Cs = torch.empty(hi - lo, K, device="cuda")
        if hi > lo:
            pkg.set_csrmm_segment(pkg.csrmm_segment(nnz, m=n))  # the WHOLE matrix's segment length on every shard
            try:
                pkg.csrmm(rps, cis, vs, B, Cs, hi - lo, n, K)
            finally:
                pkg.set_csrmm_segment(0)
        outs.append(Cs)
    assert torch.equal(torch.cat(outs), full)


def _mg_worker(rank, world, port, out):
    import torch.distributed as dist
    import spmm_denseblock_b200 as pkg
    from spmm_denseblock_b200 import shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)  # only carries the 128-byte NCCL id
    rng = np.random.default_rng(31)
    n, K = 9000, 256
    rp, ci, v = random_csr(rng, n, n, 10, hubs={7: 5000, 4100: 2500, 8999: 700}, empty_rows=range(50, 70))
    Bfull = rng.uniform(-1, 1, (n, K)).astype(np.float32)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    rpd, cid, vd, Bd = d(rp), d(ci), d(v), d(Bfull)
    full = torch.empty(n, K, device="cuda")
    pkg.csrmm(rpd, cid, vd, Bd, full, n, n, K)
    ranges = shard.shard_plan(rpd, world)
    lo, hi = ranges[rank]
    mg = shard.MultiGpu(rank, world)
    ok = True
    # all-gather-v in place
    Bg = torch.full((n, K), float("nan"), device="cuda")
    Bg[lo:hi] = Bd[lo:hi]
    mg.allgather_rows(Bg, ranges)
    torch.cuda.synchronize()
    ok &= bool(torch.equal(Bg, Bd))
    # sharded SpMM from this rank's rows of B, every slab count, twice in a row (replica reuse)
    rps, cis, vs = shard.shard_csr(rpd, cid, vd, lo, hi)
    pkg.set_csrmm_segment(pkg.csrmm_segment(ci.size, m=n))
    for slabs in (1, 2, 4, 0, 4):
        Cs = torch.full((hi - lo, K), float("nan"), device="cuda")
        mg.sharded_csrmm(rps, cis, vs, Bd[lo:hi].contiguous(), ranges, Cs, n, K, slabs=slabs)
        torch.cuda.synchronize()
        ok &= bool(torch.equal(Cs, full[lo:hi]))
    pkg.set_csrmm_segment(0)
    mg.close()
    res = [None] * world
    dist.all_gather_object(res, ok)
    if rank == 0:
        out.put(all(res))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("world", [1, 2])
def test_sharded_csrmm_c_entry_nccl(world):
    """spmmdb_mg_* + spmmdb_sharded_csrmm_f32 (C++ / NCCL): all-gather-v of B in one send / recv group, feature-slab
    pipeline; each rank's rows are the bits of the single-GPU product.  world = 2 needs two GPUs."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_mg_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    assert out.get(timeout=5) is True

"""Developer check of the tcgen05 BSR kernel against the oracle on small cases (prints errors, no asserts)."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import spmm_denseblock_b200 as pkg
from oracle import oracle as O
from helpers import random_bsr

dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
cases = []
for math in (1, 2):
    for bs in (16, 32, 64):
        for direction in (0, 1):
            for (mb, nb, n, avg) in ((3, 4, 128, 2), (67, 59, 128, 9), (40, 33, 300, 6), (20, 25, 40, 5)):
                cases.append((math, bs, direction, mb, nb, n, avg))
only = sys.argv[1:] and int(sys.argv[1])
for idx, (math, bs, direction, mb, nb, n, avg) in enumerate(cases):
    if only and idx >= only:
        break
    rng = np.random.default_rng(idx)
    brp, bci, bv = random_bsr(rng, mb, nb, bs, avg)
    k, m = nb * bs, mb * bs
    B = rng.uniform(-1, 1, (k, n)).astype(np.float32)
    C0 = rng.uniform(-1, 1, (m, n)).astype(np.float32)
    C = dev(C0)
    try:
        pkg.bsrmm(direction, 1, mb, n, nb, bci.size, 0.5, dev(bv), dev(brp), dev(bci), bs, dev(B), n, 2.0, C, n,
                  c_layout=0, math=math)
        torch.cuda.synchronize()
    except Exception as e:
        print(idx, (math, bs, direction, mb, nb, n), "EXC", e, flush=True)
        continue
    ref = O.bsrmm(brp, bci, bv, bs, B, n, n, 1, direction=direction, alpha=0.5, beta=2.0, C0=C0, ldc=n, c_layout=0)
    scale = 0.5 * O.bsrmm(brp, bci, np.abs(bv), bs, np.abs(B), n, n, 1, direction=direction) + 2 * np.abs(C0)
    got = C.cpu().numpy().astype(np.float64)
    err = np.abs(got - ref) / np.maximum(scale, 1e-30)
    bad = np.argwhere(err > 1e-2)
    print(idx, dict(math=math, bs=bs, dir=direction, mb=mb, nb=nb, n=n), "max rel err %.3e" % err.max(),
          "bad %d/%d" % (len(bad), err.size), "first bad", bad[:3].tolist(), flush=True)

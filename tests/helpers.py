"""Shared input builders for the parity tests (numpy, seeded)."""
import numpy as np


def random_csr(rng, m, n, avg_deg, hubs=(), empty_rows=(), values="uniform"):
    rows = []
    hubs = dict(hubs)
    for i in range(m):
        d = int(rng.poisson(avg_deg))
        if i in hubs:
            d = hubs[i]
        if i in empty_rows:
            d = 0
        rows.append(np.sort(rng.choice(n, size=min(d, n), replace=False)))
    rowptr = np.zeros(m + 1, np.int32)
    rowptr[1:] = np.cumsum([len(r) for r in rows])
    colind = np.concatenate(rows).astype(np.int32) if rowptr[-1] else np.zeros(0, np.int32)
    if values == "unit":
        val = np.ones(colind.size, np.float32)
    else:
        val = rng.uniform(-1, 1, colind.size).astype(np.float32)
    return rowptr, colind, val


def random_bsr(rng, mb, nb, bs, avg_blocks, values="uniform"):
    rows = []
    for _ in range(mb):
        d = min(int(rng.poisson(avg_blocks)), nb)
        rows.append(np.sort(rng.choice(nb, size=d, replace=False)))
    rowptr = np.zeros(mb + 1, np.int32)
    rowptr[1:] = np.cumsum([len(r) for r in rows])
    colind = np.concatenate(rows).astype(np.int32) if rowptr[-1] else np.zeros(0, np.int32)
    val = rng.uniform(-1, 1, colind.size * bs * bs).astype(np.float32)
    if values == "sparse":  # blocks with explicit zeros, like csr2bsr output
        val *= (rng.uniform(0, 1, val.size) < 0.3)
    return rowptr, colind, val.astype(np.float32)


def rel_err(got, ref64, scale):
    """max |got - ref| / scale, scale = per-element sum |a_ij b_jk| (>= tiny)."""
    import numpy as np
    return float(np.max(np.abs(got.astype(np.float64) - ref64) / np.maximum(scale, 1e-30)))

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


def undirected_graph(rng, n, avg_deg, components=1, isolated=(), hubs=(), self_loops=0, duplicates=0, shuffle_rows=False):
    """Symmetric adjacency pattern as CSR (rowptr, colind): `components` groups of vertices (interleaved ids, so
    that a component's ids are not contiguous) with no edge between groups; optional hub vertices (id -> degree),
    isolated vertices, self loops and duplicated entries; shuffle_rows leaves each list in random order."""
    import scipy.sparse as sp
    group = np.arange(n) % max(components, 1)
    comps = max(components, 1)
    m = int(n * avg_deg / 2)
    a = rng.integers(0, n, 2 * m + 8)
    b = rng.integers(0, (n + comps - 1) // comps, 2 * m + 8) * comps + group[a]  # a member of a's group
    keep = (b < n) & (a != b)
    a, b = a[keep][:m], b[keep][:m]
    for h, d in dict(hubs).items():
        peers = np.flatnonzero((group == group[h]) & (np.arange(n) != h))
        peers = rng.choice(peers, size=min(d, peers.size), replace=False)
        a = np.concatenate([a, np.full(peers.size, h)])
        b = np.concatenate([b, peers])
    iso = np.zeros(n, bool)
    iso[list(isolated)] = True
    keep = ~(iso[a] | iso[b])
    a, b = a[keep], b[keep]
    A = sp.coo_matrix((np.ones(a.size, np.int8), (a, b)), shape=(n, n)).tocsr()
    A = ((A + A.T) > 0).astype(np.int8).tolil()
    loops = rng.choice(np.flatnonzero(~iso), size=min(self_loops, int((~iso).sum())), replace=False) if self_loops else []
    for v in loops:
        A[v, v] = 1
    A = A.tocsr()
    A.sort_indices()
    rows = [A.indices[A.indptr[i]:A.indptr[i + 1]].astype(np.int32) for i in range(n)]
    for _ in range(duplicates):  # a neighbour listed twice (both directions keep the pattern symmetric)
        i = int(rng.integers(0, n))
        if rows[i].size:
            j = int(rows[i][rng.integers(0, rows[i].size)])
            rows[i] = np.sort(np.append(rows[i], j))
            if j != i:
                rows[j] = np.sort(np.append(rows[j], i))
    if shuffle_rows:
        rows = [rng.permutation(r) for r in rows]
    rowptr = np.zeros(n + 1, np.int32)
    rowptr[1:] = np.cumsum([r.size for r in rows])
    colind = np.concatenate(rows).astype(np.int32) if rowptr[-1] else np.zeros(0, np.int32)
    return rowptr, colind

"""O(nnz) synthetic graph generators for the BASELINE.json config shapes (the
OGB datasets are not available offline).  Pure torch index
arithmetic, so the same code runs on the CPU (tests, small sizes) and on the
GPU (bench, full sizes).  Every generator returns a CSR pattern with ascending,
duplicate-free column indices per row -- the contract the reference's loaders
establish (load_data.cc:179-182 sorts neighbours).

The reference's own generator (load_data.cc:42-69 randomCSRMatrix) draws one
Bernoulli per matrix entry, O(m*n); `bernoulli_csr` draws the same distribution
in O(nnz) by sampling the gaps.
"""
import math

import torch


def _gen(seed, device):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return g


def _csr_from_pairs(rows, cols, n_rows, n_cols, target=None, draw=None, g=None):
    """Sort (row, col) pairs, drop duplicates, build int32 CSR.  With `target` and `draw` (a function
    e -> (rows, cols) of e more edges from the same distribution) the de-duplicated edge set is topped up
    to exactly `target` entries, so a config's stated nnz is met, not undershot by the duplicates."""
    key = rows.to(torch.int64) * n_cols + cols.to(torch.int64)
    key = torch.unique(key)  # sorted + de-duplicated
    if target is not None and draw is not None:
        for _ in range(12):
            short = target - key.numel()
            if short <= 0:
                break
            r2, c2 = draw(int(short * 1.05) + 1024)
            key = torch.unique(torch.cat([key, r2.to(torch.int64) * n_cols + c2.to(torch.int64)]))
        if key.numel() > target:  # drop the surplus uniformly
            keep = torch.randperm(key.numel(), generator=g, device=key.device)[:target]
            key = key[torch.sort(keep).values]
    r = torch.div(key, n_cols, rounding_mode="floor")
    c = (key - r * n_cols).to(torch.int32)
    counts = torch.bincount(r, minlength=n_rows)
    rowptr = torch.zeros(n_rows + 1, dtype=torch.int64, device=key.device)
    rowptr[1:] = torch.cumsum(counts, 0)
    return rowptr.to(torch.int32), c


def uniform_degree_graph(n, degree, seed=1234, device="cpu"):
    """C1: every row draws `degree` distinct uniform columns (sorted)."""
    g = _gen(seed, device)
    rows = torch.arange(n, device=device).repeat_interleave(degree)
    # distinct columns per row: stratified draw (one per n/degree stripe) keeps it O(nnz)
    stripe = n // degree
    base = torch.arange(degree, device=device).repeat(n) * stripe
    cols = base + torch.randint(0, stripe, (n * degree,), generator=g, device=device)
    return _csr_from_pairs(rows, cols, n, n)


def bernoulli_csr(m, n, p, seed=1234, device="cpu"):
    """Each entry non-zero with probability p (load_data.cc:42-69), in O(nnz):
    positions of successes in the flattened m*n Bernoulli stream via geometric gaps."""
    g = _gen(seed, device)
    total = m * n
    expect = int(total * p)
    draw = int(expect + 6 * math.sqrt(max(expect, 1)) + 64)
    u = torch.rand(draw, generator=g, device=device, dtype=torch.float64).clamp_(min=1e-300)
    gaps = torch.floor(torch.log(u) / math.log1p(-p)).to(torch.int64) + 1
    pos = torch.cumsum(gaps, 0) - 1
    pos = pos[pos < total]
    rows = torch.div(pos, n, rounding_mode="floor")
    cols = pos - rows * n
    return _csr_from_pairs(rows, cols, m, n)


def power_law_degrees(n, nnz, max_degree, gamma=2.2, seed=1234, device="cpu"):
    """Zipf rank-size degrees (exponent 1/(gamma-1)), capped, rescaled to sum ~ nnz, shuffled."""
    g = _gen(seed, device)
    rank = torch.arange(1, n + 1, device=device, dtype=torch.float64)
    w = rank.pow(-1.0 / (gamma - 1.0))
    lo, hi = 0.0, float(nnz)
    for _ in range(60):  # scale so that sum(clamp(c*w, 1, max_degree)) == nnz
        c = 0.5 * (lo + hi)
        s = torch.clamp(c * w, 1.0, float(max_degree)).sum().item()
        if s < nnz:
            lo = c
        else:
            hi = c
    d = torch.clamp(hi * w, 1.0, float(max_degree))
    deg = torch.floor(d).to(torch.int64)
    frac = d - deg
    deg += (torch.rand(n, generator=g, device=device, dtype=torch.float64) < frac).to(torch.int64)
    perm = torch.randperm(n, generator=g, device=device)
    return deg[perm]


def power_law_graph(n, nnz, max_degree, gamma=2.2, p_local=0.0, window=2048, seed=1234, device="cpu", exact=True):
    """C2 / C5: power-law out-degrees; each edge's column is, with probability
    p_local, a neighbour id within ~`window` of the row (the locality a
    community-aware reordering such as Rabbit / Gorder leaves behind), otherwise
    drawn preferentially (probability proportional to degree: hubs)."""
    g = _gen(seed + 1, device)
    deg = power_law_degrees(n, nnz, max_degree, gamma, seed, device)
    rows = torch.arange(n, device=device).repeat_interleave(deg)
    cdf = torch.cumsum(deg.to(torch.float64), 0)

    def columns(rows):
        e = rows.numel()
        u = torch.rand(e, generator=g, device=device, dtype=torch.float64) * cdf[-1]
        cols = torch.searchsorted(cdf, u).clamp_(max=n - 1)
        if p_local > 0:
            local = torch.rand(e, generator=g, device=device) < p_local
            off = torch.round(torch.randn(e, generator=g, device=device) * (window / 2.0)).to(torch.int64)
            cols = torch.where(local, torch.remainder(rows + off, n), cols)
        return cols

    def draw(e):  # e more edges: source rows in proportion to their degree, columns as above
        u = torch.rand(e, generator=g, device=device, dtype=torch.float64) * cdf[-1]
        r = torch.searchsorted(cdf, u).clamp_(max=n - 1)
        return r, columns(r)

    return _csr_from_pairs(rows, columns(rows), n, n, target=nnz if exact else None, draw=draw, g=g)


def planted_community_graph(n, nnz, min_size=256, max_size=2048, p_in=0.9, shuffle=False, seed=1234,
                            device="cpu", exact=True):
    """C3: stochastic-block-model-like graph: contiguous communities of random size,
    a fraction p_in of every row's edges stays inside its community.  With
    shuffle=True node ids are randomly permuted (what a reordering has to undo);
    returns (rowptr, colind, perm) where perm[old] = new for the shuffle (or None)."""
    g = _gen(seed, device)
    sizes = []
    left = n
    cpu_g = torch.Generator().manual_seed(seed)
    while left > 0:
        s = int(torch.randint(min_size, max_size + 1, (1,), generator=cpu_g).item())
        s = min(s, left)
        sizes.append(s)
        left -= s
    sizes = torch.tensor(sizes, device=device)
    starts = torch.cumsum(sizes, 0) - sizes
    comm = torch.arange(sizes.numel(), device=device).repeat_interleave(sizes)
    deg = max(1, nnz // n)
    rows = torch.arange(n, device=device).repeat_interleave(deg)
    perm = torch.randperm(n, generator=g, device=device) if shuffle else None

    def columns(rows):
        e = rows.numel()
        inside = torch.rand(e, generator=g, device=device) < p_in
        u = torch.rand(e, generator=g, device=device, dtype=torch.float64)
        c_in = starts[comm[rows]] + torch.floor(u * sizes[comm[rows]].to(torch.float64)).to(torch.int64)
        c_out = torch.floor(u * n).to(torch.int64)
        cols = torch.where(inside, c_in, c_out).clamp_(max=n - 1)
        return (perm[rows], perm[cols]) if shuffle else (rows, cols)

    def draw(e):
        return columns(torch.randint(0, n, (e,), generator=g, device=device))

    r0, c0 = columns(rows)
    rp, ci = _csr_from_pairs(r0, c0, n, n, target=nnz if exact else None, draw=draw, g=g)
    return rp, ci, perm


def molecule_batch_graph(num_graphs, mean_nodes=25.5, min_nodes=2, max_nodes=222, extra_edges=3, seed=1234,
                         device="cpu"):
    """C4: block-diagonal batch of small undirected graphs (ogbg_molhiv.py:79-106 shape):
    each graph is a random tree plus `extra_edges` ring closures, both directions stored."""
    g = _gen(seed, device)
    nodes = torch.round(torch.randn(num_graphs, generator=g, device=device) * 12.0 + mean_nodes)
    nodes = nodes.clamp_(min_nodes, max_nodes).to(torch.int64)
    starts = torch.cumsum(nodes, 0) - nodes
    n = int(nodes.sum().item())
    gid = torch.arange(num_graphs, device=device).repeat_interleave(nodes)
    local = torch.arange(n, device=device) - starts[gid]
    # tree: every node but the first of its graph picks a parent among earlier nodes
    child = torch.nonzero(local > 0).squeeze(1)
    u = torch.rand(child.numel(), generator=g, device=device, dtype=torch.float64)
    parent = starts[gid[child]] + torch.floor(u * local[child].to(torch.float64)).to(torch.int64)
    # ring closures: random pairs inside a graph
    gsel = torch.arange(num_graphs, device=device).repeat_interleave(extra_edges)
    ua = torch.rand(gsel.numel(), generator=g, device=device, dtype=torch.float64)
    ub = torch.rand(gsel.numel(), generator=g, device=device, dtype=torch.float64)
    a = starts[gsel] + torch.floor(ua * nodes[gsel].to(torch.float64)).to(torch.int64)
    b = starts[gsel] + torch.floor(ub * nodes[gsel].to(torch.float64)).to(torch.int64)
    keep = a != b
    src = torch.cat([child, parent, a[keep], b[keep]])
    dst = torch.cat([parent, child, b[keep], a[keep]])
    rp, ci = _csr_from_pairs(src, dst, n, n)
    return rp, ci, nodes


def apply_permutation(rowptr, colind, old2new):
    """A' = P A P^T on a pattern-only CSR (reorder_strategy.cc:26-39 permutate), torch version
    used by the generators/bench; the bit-exactness contract is checked in tests/."""
    n = rowptr.numel() - 1
    deg = (rowptr[1:] - rowptr[:-1]).to(torch.int64)
    rows = torch.arange(n, device=rowptr.device).repeat_interleave(deg)
    key = old2new[rows].to(torch.int64) * n + old2new[colind.to(torch.int64)].to(torch.int64)
    key, _ = torch.sort(key)
    r = torch.div(key, n, rounding_mode="floor")
    counts = torch.bincount(r, minlength=n)
    rp = torch.zeros(n + 1, dtype=torch.int64, device=rowptr.device)
    rp[1:] = torch.cumsum(counts, 0)
    return rp.to(torch.int32), (key - r * n).to(torch.int32)


CONFIGS = {
    # name: (generator kwargs, K) -- BASELINE.json configs[0..4]
    "c1_uniform": dict(kind="uniform", n=65536, degree=16, K=64),
    "c2_arxiv": dict(kind="power_law", n=169343, nnz=1166243, max_degree=13000, K=128),
    "c3_proteins": dict(kind="community", n=132534, nnz=39561252, K=128),
    "c4_molhiv": dict(kind="molecules", num_graphs=41127, K=300),
    "c5_products": dict(kind="power_law", n=2449029, nnz=61859140, max_degree=17500, K=256),
}


def make_config_graph(name, device="cpu", seed=1234, scale=1.0, **overrides):
    """Build the pattern for a BASELINE.json config; `scale` < 1 shrinks node and edge counts
    proportionally (parity tests run the same generator at small sizes)."""
    cfg = dict(CONFIGS[name])
    cfg.update(overrides)
    kind = cfg["kind"]
    if kind == "uniform":
        n = max(64, int(cfg["n"] * scale))
        return uniform_degree_graph(n, cfg["degree"], seed, device)
    if kind == "power_law":
        n = max(64, int(cfg["n"] * scale))
        nnz = max(n, int(cfg["nnz"] * scale))
        md = max(8, min(cfg["max_degree"], n // 2))
        return power_law_graph(n, nnz, md, p_local=cfg.get("p_local", 0.0), window=cfg.get("window", 2048),
                               seed=seed, device=device, exact=cfg.get("exact", True))
    if kind == "community":
        n = max(1024, int(cfg["n"] * scale))
        nnz = max(n, int(cfg["nnz"] * scale))
        rp, ci, _ = planted_community_graph(n, nnz, shuffle=cfg.get("shuffle", False), seed=seed, device=device,
                                            min_size=cfg.get("min_size", 256), max_size=cfg.get("max_size", 2048),
                                            p_in=cfg.get("p_in", 0.9), exact=cfg.get("exact", True))
        return rp, ci
    if kind == "molecules":
        rp, ci, _ = molecule_batch_graph(max(2, int(cfg["num_graphs"] * scale)), seed=seed, device=device)
        return rp, ci
    raise ValueError(kind)

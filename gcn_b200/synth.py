"""Synthetic graphs of the shapes BASELINE.json names: there is no network for the
real datasets, so every measurement runs on seeded random graphs with the same vertex / edge
counts.  Generation happens with torch on whatever device is given (the GPU for the large shapes —
it is input preparation, not the measured path) and returns the normalised adjacency
Â = D^-½ (A + I) D^-½ as a host CSR (int32 rowptr/col, fp32 values, ascending columns): what
utils.normalize_adj + to_sparse_csr produce in the reference (pygcn/gcnio/util/utils.py:78-90,
gcn6.py:302-312).  Values are computed in fp64 and cast to fp32 like the reference does.

Shapes (undirected edge counts before the self-loops are added):
  cora      n=2 708        5 278 edges    Erdős–Rényi pairs
  arxiv     n=169 343      1 166 243      random pairs, symmetrised
  reddit    n=232 965      57.3 M         Chung-Lu power law (exponent 2.1), 57.3 M DISTINCT pairs
                                          (nnz(Â) = 2·57.3 M + n = 114.8 M, the "114 M edges" of the shape)
  products  n=2 449 029    61.86 M        power-law degrees, community-local endpoints (see
                                          `local_powerlaw`): co-purchase graphs are strongly
                                          clustered; ids are community-contiguous
  rmat      n=2^scale      ef·n           R-MAT (a,b,c,d) = (.57,.19,.19,.05), Graph500 style
"""
from __future__ import annotations

import numpy as np
import torch

SEED = 15  # the reference's seed (profiling_gcn.py:76-80)

SHAPES = {
    "cora": dict(n=2708, edges=5278, kind="er"),
    "arxiv": dict(n=169_343, edges=1_166_243, kind="er"),
    "reddit": dict(n=232_965, edges=57_300_000, kind="chung_lu", exponent=2.1),
    "products": dict(n=2_449_029, edges=61_859_140, kind="local", exponent=2.6, p_local=0.9,
                     window=2048),
}


def _gen(device, seed):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return g


def _powerlaw_cdf(n, exponent, device, g):
    # Chung-Lu weights w_i ∝ (i+1)^(-1/(γ-1)), assigned to vertices in random order
    w = torch.arange(1, n + 1, device=device, dtype=torch.float64).pow_(-1.0 / (exponent - 1.0))
    w = w[torch.randperm(n, device=device, generator=g)]
    cdf = torch.cumsum(w, 0)
    return cdf / cdf[-1]


def _sample(cdf, count, g):
    u = torch.rand(count, device=cdf.device, dtype=torch.float64, generator=g)
    return torch.searchsorted(cdf, u).clamp_(max=cdf.numel() - 1)


def edges_er(n, edges, device, seed=SEED):
    g = _gen(device, seed)
    return (torch.randint(0, n, (edges,), device=device, generator=g),
            torch.randint(0, n, (edges,), device=device, generator=g))


def edges_chung_lu(n, edges, exponent, device, seed=SEED):
    """`edges` DISTINCT undirected non-loop pairs with Chung-Lu (power-law weight) endpoints.

    Sampling with replacement loses a large share of the pairs as duplicates on a hub-heavy sequence
    (exponent 2.1, n = 233 k: 57.3 M samples leave 34 M distinct pairs, i.e. a graph 40 % smaller than
    the shape names — round-1 VERDICT item 4), so pairs are drawn in rounds until enough distinct ones
    exist and the surplus is dropped uniformly at random."""
    g = _gen(device, seed)
    cdf = _powerlaw_cdf(n, exponent, device, g)
    keys = torch.empty(0, dtype=torch.int64, device=device)
    draw = edges
    for _round in range(64):
        a, b = _sample(cdf, draw, g), _sample(cdf, draw, g)
        lo, hi = torch.minimum(a, b), torch.maximum(a, b)
        keep = lo != hi
        keys = torch.unique(torch.cat((keys, lo[keep] * n + hi[keep])))
        del a, b, lo, hi, keep
        if keys.numel() >= edges:
            break
        # the yield of a round falls as the hub pairs fill up: draw the shortfall times the inverse of
        # the last yield (bounded), so that a handful of rounds suffices
        draw = int(min(4 * edges, max(edges // 8, 2.5 * (edges - keys.numel()))))
    if keys.numel() < edges:
        raise RuntimeError(f"chung_lu: only {keys.numel()} distinct pairs of {edges} after 64 rounds")
    if keys.numel() > edges:
        pick = torch.randperm(keys.numel(), device=device, generator=g)[:edges]
        keys = keys[pick]
    src = torch.div(keys, n, rounding_mode="floor")
    return src, keys - src * n


def edges_local_powerlaw(n, edges, exponent, p_local, window, device, seed=SEED):
    """Sources by power-law weight; a fraction p_local of the destinations within a two-sided
    geometric window around the source id (community-contiguous numbering), the rest by power-law
    weight anywhere (hubs)."""
    g = _gen(device, seed)
    cdf = _powerlaw_cdf(n, exponent, device, g)
    src = _sample(cdf, edges, g)
    far = _sample(cdf, edges, g)
    mag = torch.empty(edges, device=device, dtype=torch.float64).exponential_(1.0 / window, generator=g)
    sign = torch.randint(0, 2, (edges,), device=device, generator=g) * 2 - 1
    near = (src + sign * (mag.long() + 1)).remainder_(n)
    pick = torch.rand(edges, device=device, generator=g) < p_local
    return src, torch.where(pick, near, far)


def edges_rmat(scale, edge_factor, device, abcd=(0.57, 0.19, 0.19, 0.05), seed=SEED, chunk=1 << 26):
    n = 1 << scale
    total = edge_factor * n
    g = _gen(device, seed)
    a, b, c, _ = abcd
    srcs, dsts = [], []
    for start in range(0, total, chunk):
        cnt = min(chunk, total - start)
        src = torch.zeros(cnt, dtype=torch.int64, device=device)
        dst = torch.zeros(cnt, dtype=torch.int64, device=device)
        for _level in range(scale):
            r = torch.rand(cnt, device=device, generator=g)
            sbit = (r >= a + b).long()                       # quadrants c, d: lower half of rows
            dbit = (((r >= a) & (r < a + b)) | (r >= a + b + c)).long()   # quadrants b, d
            src = (src << 1) | sbit
            dst = (dst << 1) | dbit
        srcs.append(src)
        dsts.append(dst)
    return torch.cat(srcs), torch.cat(dsts)


def _unique_sorted(key, n):
    """torch.unique(key) (sorted), in row-range pieces when key has 2^31 or more elements (the CUB sort behind
    torch.unique takes at most 2^31 - 1: R-MAT scale 24 with edge factor 64 has 2.16 G keys)."""
    if key.numel() < 2 ** 31 - 1:
        return torch.unique(key)
    pieces = int(key.numel() // (2 ** 30)) + 1
    parts = []
    for i in range(pieces):
        lo, hi = (n * i) // pieces, (n * (i + 1)) // pieces
        sel = key[(key >= lo * n) & (key < hi * n)]
        parts.append(torch.unique(sel))
        del sel
    return torch.cat(parts)


def normalized_csr(src, dst, n, to_host=True):
    """Symmetrise, drop duplicates and the diagonal, add I, scale by D^-½ on both sides.
    Returns (rowptr int32, col int32, val fp32) as numpy arrays (or device tensors)."""
    dev = src.device
    diag = torch.arange(n, device=dev, dtype=torch.int64)
    keep = src != dst
    src, dst = src[keep], dst[keep]
    key = torch.cat((src * n + dst, dst * n + src, diag * n + diag))
    del src, dst, keep
    key = _unique_sorted(key, n)  # sorted: row-major, ascending columns
    nnz = key.numel()
    if nnz >= 2 ** 31 - 16:
        raise ValueError(f"{nnz} non-zeros do not fit the int32 CSR offsets of one plan "
                         "(include/gcn_b200.h: nnz < 2^31 - 16); partition the rows first")
    # row pointers by binary search of the row boundaries in the sorted keys (no 2^31 limits, no histogram)
    rowptr = torch.searchsorted(key, diag * n)
    rowptr = torch.cat((rowptr, torch.tensor([nnz], device=dev, dtype=rowptr.dtype)))
    counts = rowptr[1:] - rowptr[:-1]
    dinv = counts.to(torch.float64).pow_(-0.5)   # row sum of (A + I) with unit weights = its nnz
    col = torch.empty(nnz, dtype=torch.int32, device=dev)
    val = torch.empty(nnz, dtype=torch.float32, device=dev)
    step = 1 << 28
    for z0 in range(0, nnz, step):               # in pieces: the fp64 temporaries of 2 G non-zeros would take 50 GB
        k = key[z0:z0 + step]
        r = torch.div(k, n, rounding_mode="floor")
        c = k - r * n
        col[z0:z0 + step] = c.to(torch.int32)
        val[z0:z0 + step] = (dinv[r] * dinv[c]).to(torch.float32)
    del key
    rowptr = rowptr.to(torch.int32)
    if not to_host:
        return rowptr, col, val
    return rowptr.cpu().numpy(), col.cpu().numpy(), val.cpu().numpy()


def make_graph(name, device="cpu", seed=SEED, scale=None, edge_factor=None, shuffle=False, **over):
    """(rowptr, col, val, n) of Â for a named shape; `rmat` takes scale / edge_factor.
    shuffle=True relabels the vertices by a random permutation (hides any id locality, the
    starting point of the reorder sweep)."""
    device = torch.device(device)
    if name == "rmat":
        n = 1 << scale
        src, dst = edges_rmat(scale, edge_factor, device, seed=seed)
    else:
        spec = dict(SHAPES[name])
        spec.update(over)
        n = spec["n"]
        if spec["kind"] == "er":
            src, dst = edges_er(n, spec["edges"], device, seed)
        elif spec["kind"] == "chung_lu":
            src, dst = edges_chung_lu(n, spec["edges"], spec["exponent"], device, seed)
        else:
            src, dst = edges_local_powerlaw(n, spec["edges"], spec["exponent"], spec["p_local"],
                                            spec["window"], device, seed)
    if shuffle:
        perm = torch.randperm(n, device=device, generator=_gen(device, seed + 1))
        src, dst = perm[src], perm[dst]
    rowptr, col, val = normalized_csr(src, dst, n)
    return rowptr, col, val, n


def spmm_bytes_flops(m, n, nnz, k, b_bytes=4, c_bytes=4):
    """Algorithmic bytes and flops of one SpMM: CSR streamed once
    (int32 col + fp32 val + int32 rowptr), B read once, C written once."""
    return 8 * nnz + 4 * (m + 1) + b_bytes * n * k + c_bytes * m * k, 2 * nnz * k

#!/usr/bin/env python
"""Generates tests/golden/*.npz by EXECUTING THE REFERENCE in the build container.

Run from the repo root, once, where /root/reference is mounted:
    make -C oracle            # builds oracle/_ref/*.so from the reference sources where they lie
    python tests/golden/make_golden.py

Sources of every stored answer (nothing here is computed by gcn_b200 or by oracle/'s restatements):
  * Â and the COO tensor: the reference's Python, imported from /root/reference —
    utils.normalize_adj (pygcn/gcnio/util/utils.py:78-90) and
    utils.sparse_mx_to_torch_sparse_tensor (:243-250);
  * SpMM answers: torch.spmm(adj, B) on CPU — the call the reference layer makes
    (pygcn/gcn1.py:53, gcn5.py:90) — plus scipy CSR @ in fp64 as the independent exact value;
  * orders / renumbering: the compiled, unmodified reference host code (oracle/_ref/ref_graph.so:
    order_deg.cu, order_rcm.cu, order_gorder.cu, renumber.cu dfs/gorder/perm_apply);
  * tile metadata: compiled tile.cu, as shipped (n_segs, segVoMap) and with the one-token
    tile.cu:71 correction (seg_rowPtr, segNzCV) — SURVEY.md §8c;
  * GCN: pygcn.gcn1.GCN forward + backward on CPU with dropout disabled (Cora-shaped);
  * (Â·X)·W + b layer: pygcn.gcn5.GraphConvolution2 forward + backward on CPU (gc2_powerlaw.npz).
The fixtures travel to the GPU box; /root/reference does not.
"""
import os
import sys

import numpy as np
import scipy.sparse as sp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from oracle import ref_bindings as R  # noqa: E402
from pygcn.gcnio.util import utils as ref_utils  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
SEED = 15  # profiling_gcn.py:76-80


def pattern(kind, n, e, rng):
    if kind == "er":
        r, c = rng.integers(0, n, e), rng.integers(0, n, e)
    elif kind == "powerlaw":
        w = (np.arange(1, n + 1, dtype=np.float64)) ** (-1.0 / 1.1)
        w = rng.permutation(w)
        cdf = np.cumsum(w) / w.sum()
        r = np.searchsorted(cdf, rng.random(e)).clip(max=n - 1)
        c = np.searchsorted(cdf, rng.random(e)).clip(max=n - 1)
    elif kind == "banded":
        r = rng.integers(0, n, e)
        c = (r + rng.integers(-6, 7, e)) % n
    elif kind == "star":  # one row far above the 116-nnz segment limit and the 512 block class
        r = np.concatenate((np.zeros(e, dtype=np.int64), rng.integers(0, n, e // 4)))
        c = np.concatenate((rng.integers(0, n, e), rng.integers(0, n, e // 4)))
    a = sp.coo_matrix((np.ones(len(r)), (r, c)), shape=(n, n)).tocsr()
    a = ((a + a.T) > 0).astype(np.float64)
    a.setdiag(0)
    a.eliminate_zeros()
    return a


def ref_normalized_csr(a):
    mx = ref_utils.normalize_adj(a)                               # fp64 scipy
    coo = ref_utils.sparse_mx_to_torch_sparse_tensor(mx).coalesce()  # fp32 COO, the layer's operand
    csr = sp.csr_matrix(mx).astype(np.float32)
    csr.sort_indices()
    return mx, coo, csr


def graph_case(name, kind, n, e, k, seed):
    rng = np.random.default_rng(seed)
    a = pattern(kind, n, e, rng)
    mx64, coo, csr = ref_normalized_csr(a)
    rp, ci, va = csr.indptr.astype(np.int32), csr.indices.astype(np.int32), csr.data
    B = rng.standard_normal((n, k)).astype(np.float32)
    out = dict(a_rowptr=a.indptr.astype(np.int32), a_col=a.indices.astype(np.int32),
               rowptr=rp, col=ci, val=va, B=B)
    out["C_torch_spmm"] = torch.spmm(coo, torch.from_numpy(B)).numpy()
    out["C_f64"] = (sp.csr_matrix((va.astype(np.float64), ci, rp), shape=(n, n)) @ B.astype(np.float64))
    # unsymmetric operand for the transposed product: keep the strict upper triangle + diagonal
    up = sp.triu(csr).tocsr()
    up.sort_indices()
    out["U_rowptr"], out["U_col"], out["U_val"] = up.indptr.astype(np.int32), up.indices.astype(np.int32), up.data
    ucoo = ref_utils.sparse_mx_to_torch_sparse_tensor(up).coalesce()
    Bt = torch.from_numpy(B).requires_grad_(True)
    G = rng.standard_normal((n, k)).astype(np.float32)
    torch.spmm(ucoo, Bt).backward(torch.from_numpy(G))           # autograd's Aᵀ·dY
    out["G"] = G
    out["UtG_autograd"] = Bt.grad.numpy()
    # host preprocessing by the compiled reference
    out["rank_deg_desc"] = R.ref_order_deg(rp, ci, True)
    out["rank_deg_asc"] = R.ref_order_deg(rp, ci, False)
    out["rank_rcm"] = R.ref_order_rcm(rp, ci)
    out["rank_gorder3"] = R.ref_order_gorder(rp, ci, 3)
    out["rank_gorder5"] = R.ref_order_gorder(rp, ci, 5)
    for nm, fn in (("gorder", R.ref_gorder), ("dfs", R.ref_dfs), ("rabbit", R.ref_rabbit)):
        r2, c2, v2, vomp = fn(rp, ci, va)
        out[f"{nm}_rowptr"], out[f"{nm}_col"], out[f"{nm}_val"], out[f"{nm}_vomp"] = r2, c2, v2, vomp
    perm = rng.permutation(n).astype(np.int32)
    r2, c2, v2 = R.ref_perm_apply(rp, ci, va, perm)
    out["perm"], out["perm_rowptr"], out["perm_col"], out["perm_val"] = perm, r2, c2, v2
    shipped = R.ref_csr2seg(rp, ci, va, corrected=False)
    fixed = R.ref_csr2seg(rp, ci, va, corrected=True)
    assert shipped["n_segs"] == fixed["n_segs"] and np.array_equal(shipped["segVoMap"], fixed["segVoMap"])
    out["tile_n_segs"] = np.int64(shipped["n_segs"])
    out["tile_segVoMap"] = shipped["segVoMap"]
    out["tile_seg_rowPtr_shipped"] = shipped["seg_rowPtr"]
    out["tile_seg_rowPtr"] = fixed["seg_rowPtr"]
    out["tile_segNzCV"] = fixed["segNzCV"]
    np.savez_compressed(os.path.join(OUT, f"graph_{name}.npz"), **out)
    print(name, "n", n, "nnz", len(ci), "segs", shipped["n_segs"], "max row", int(np.diff(rp).max()))


def gcn_case():
    """BASELINE config 0: pygcn 2-layer GCN fwd+bwd, Cora-shaped (2708 nodes, 5278 undirected
    edges, 1433 -> 16 -> 7), CPU torch.spmm — pygcn/gcn1.py."""
    from pygcn.gcn1 import GCN
    rng = np.random.default_rng(SEED)
    n, nfeat, nhid, nclass = 2708, 1433, 16, 7
    a = pattern("er", n, 5278, rng)
    feats = sp.random(n, nfeat, density=0.0127, format="csr", dtype=np.float32,
                      random_state=np.random.RandomState(SEED))
    feats.data[:] = 1.0
    labels = rng.integers(0, nclass, n)
    idx_train = np.arange(int(0.05 * n))
    torch.manual_seed(SEED)
    model = GCN(nfeat=nfeat, nhid=nhid, nclass=nclass, dropout=0.0, device="cpu")
    adj_t, feat_t, lab_t = ref_utils.to_tensor(a, feats, labels, device="cpu")
    adj_norm = ref_utils.normalize_adj_tensor(adj_t, sparse=True)
    model.train()
    out = model.forward(feat_t, adj_norm)[0]
    loss = torch.nn.functional.nll_loss(out[idx_train], lab_t[idx_train])
    loss.backward()
    an = adj_norm.coalesce()
    f = feats.tocsr()
    f.sort_indices()
    np.savez_compressed(
        os.path.join(OUT, "gcn_cora.npz"),
        a_rowptr=a.indptr.astype(np.int32), a_col=a.indices.astype(np.int32),
        adj_norm_idx=an.indices().numpy(), adj_norm_val=an.values().numpy(),
        f_rowptr=f.indptr.astype(np.int32), f_col=f.indices.astype(np.int32), f_val=f.data,
        labels=labels, idx_train=idx_train,
        W1=model.gc1.weight.detach().numpy(), b1=model.gc1.bias.detach().numpy(),
        W2=model.gc2.weight.detach().numpy(), b2=model.gc2.bias.detach().numpy(),
        out=out.detach().numpy(), loss=np.float64(loss.item()),
        gW1=model.gc1.weight.grad.numpy(), gb1=model.gc1.bias.grad.numpy(),
        gW2=model.gc2.weight.grad.numpy(), gb2=model.gc2.bias.grad.numpy())
    print("gcn_cora loss", loss.item())


def gc2_case():
    """(Â·X)·W + b forward and backward by EXECUTING the reference's own layer class
    pygcn.gcn5.GraphConvolution2 (pygcn/gcn5.py:101-139: torch.spmm(adj, input), torch.mm(support, W),
    + bias) on CPU.  gcn5 loads './cuspmm.so' relative to the working directory at import
    (gcn5.py:18), so the import happens inside oracle/_ref/, where our recipe built that file."""
    cwd = os.getcwd()
    os.chdir(os.path.join(ROOT, "oracle", "_ref"))
    try:
        from pygcn import gcn5
    finally:
        os.chdir(cwd)
    rng = np.random.default_rng(SEED + 2)
    n, nin, nout = 777, 52, 24          # ragged sizes: n % 8 != 0, widths that are multiples of 4 only
    a = pattern("powerlaw", n, 6000, rng)
    adj = ref_utils.sparse_mx_to_torch_sparse_tensor(ref_utils.normalize_adj(a)).coalesce()
    torch.manual_seed(SEED)
    layer = gcn5.GraphConvolution2(nin, nout)
    x = torch.from_numpy(rng.standard_normal((n, nin)).astype(np.float32)).requires_grad_(True)
    out = layer(x, adj)
    g = torch.from_numpy(rng.standard_normal((n, nout)).astype(np.float32))
    out.backward(g)
    an = adj.coalesce()
    np.savez_compressed(
        os.path.join(OUT, "gc2_powerlaw.npz"),
        adj_idx=an.indices().numpy(), adj_val=an.values().numpy(), n=np.int64(n),
        x=x.detach().numpy(), W=layer.weight.detach().numpy(), b=layer.bias.detach().numpy(),
        out=out.detach().numpy(), gout=g.numpy(), gx=x.grad.numpy(),
        gW=layer.weight.grad.numpy(), gb=layer.bias.grad.numpy())
    print("gc2_powerlaw out", float(out.abs().max()))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "gc2":     # only the fixture added in round 2
        gc2_case()
        sys.exit(0)
    graph_case("er_small", "er", 96, 300, 16, 1)
    graph_case("er_ragged", "er", 203, 700, 7, 2)       # m % 8 != 0, odd k
    graph_case("powerlaw", "powerlaw", 400, 4000, 32, 3)
    graph_case("banded", "banded", 512, 3000, 100, 4)    # k = 100 (products width)
    graph_case("star", "star", 1500, 1300, 47, 5)        # a 1300-nnz row: block class + split segments
    gcn_case()
    gc2_case()

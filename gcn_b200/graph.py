"""Host-side graph preprocessing: vertex orders, CSR renumbering, reference tile format and the
adjacency normalisation.  numpy in / numpy out; all work happens in libgcn_b200.so
(gcn_b200/csrc/graph/*.cpp), bit-exact with the reference's host code.

Conventions are the reference's: ``order_*`` return ``rank[old] = new`` (order_rcm.cu:28-31); the
``renumber_*`` functions return ``vomp[new] = old`` (renumber.cu:184-189) together with the
renumbered CSR ``P·A·Pᵀ`` whose rows have ascending columns.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _ffi
from ._ffi import np_ptr


def _csr(rowptr, col, val=None, copy=False):
    rp = np.array(rowptr, dtype=np.int32, copy=True) if copy else np.ascontiguousarray(rowptr, dtype=np.int32)
    ci = np.array(col, dtype=np.int32, copy=True) if copy else np.ascontiguousarray(col, dtype=np.int32)
    if val is None:
        return rp, ci
    va = np.array(val, dtype=np.float32, copy=True) if copy else np.ascontiguousarray(val, dtype=np.float32)
    if len(va) != len(ci):
        raise ValueError("col and val differ in length")
    return rp, ci, va


def order_deg(rowptr, col, desc=True):
    """Rank by in+out degree, ties by vertex id (order_deg.cu:41-45)."""
    rp, ci = _csr(rowptr, col)
    m = len(rp) - 1
    out = np.empty(m, dtype=np.int64)
    _ffi.call("gcnb_order_deg", np_ptr(rp), np_ptr(ci), m, len(ci), int(bool(desc)), np_ptr(out))
    return out


def order_rcm(rowptr, col):
    """Reverse Cuthill-McKee as the reference composes it (order_rcm.cu:15-33)."""
    rp, ci = _csr(rowptr, col)
    m = len(rp) - 1
    out = np.empty(m, dtype=np.int64)
    _ffi.call("gcnb_order_rcm", np_ptr(rp), np_ptr(ci), m, len(ci), np_ptr(out))
    return out


def order_gorder(rowptr, col, window=3):
    """RCM followed by Gorder (complete_gorder, order_gorder.cu:13-31)."""
    rp, ci = _csr(rowptr, col)
    m = len(rp) - 1
    out = np.empty(m, dtype=np.int64)
    _ffi.call("gcnb_order_gorder", np_ptr(rp), np_ptr(ci), m, len(ci), int(window), np_ptr(out))
    return out


def perm_apply(rowptr, col, val, vomp):
    """P·A·Pᵀ for ``vomp[new] = old`` (renumber.cu:233-318).  Returns new (rowptr, col, val)."""
    rp, ci, va = _csr(rowptr, col, val, copy=True)
    vo = np.ascontiguousarray(vomp, dtype=np.int32)
    m = len(rp) - 1
    if len(vo) != m:
        raise ValueError("vomp length differs from the number of rows")
    _ffi.call("gcnb_perm_apply", np_ptr(rp), np_ptr(ci), np_ptr(va), np_ptr(vo), m, len(ci))
    return rp, ci, va


def _renumber(fn, rowptr, col, val, *extra):
    rp, ci, va = _csr(rowptr, col, val, copy=True)
    m = len(rp) - 1
    vomp = np.empty(m, dtype=np.int32)
    _ffi.call(fn, np_ptr(rp), np_ptr(ci), np_ptr(va), np_ptr(vomp), m, len(ci), *extra)
    return rp, ci, va, vomp


def renumber_gorder(rowptr, col, val, window=3):
    """gorder() of the reference (renumber.cu:157-230): (rowptr', col', val', vomp)."""
    return _renumber("gcnb_renumber_gorder", rowptr, col, val, int(window))


def renumber_dfs(rowptr, col, val):
    """dfs() of the reference (renumber.cu:23-155)."""
    return _renumber("gcnb_renumber_dfs", rowptr, col, val)


def renumber_rabbit(rowptr, col, val):
    """rabbit() of the reference (renumber.cu:319-522), its default order (RBT, gcn6.py:27-30)."""
    return _renumber("gcnb_renumber_rabbit", rowptr, col, val)


def renumber_rcm(rowptr, col, val):
    return _renumber("gcnb_renumber_rcm", rowptr, col, val)


def renumber_deg(rowptr, col, val, desc=True):
    return _renumber("gcnb_renumber_deg", rowptr, col, val, int(bool(desc)))


def csr2tile(rowptr, col, val, n_sm=148, tm=8):
    """CSR -> reference tile format (tile.cu:11-169) with cumulative seg_rowPtr.  n_sm replaces the
    reference's device query (148 SMs on B200).  Returns a dict of exactly-sized arrays."""
    rp, ci, va = _csr(rowptr, col, val)
    m, nnz = len(rp) - 1, len(ci)
    cnt = ctypes.c_int64(0)
    _ffi.call("gcnb_tile_count", np_ptr(rp), np_ptr(ci), m, nnz, tm, ctypes.byref(cnt))
    ns = cnt.value
    seg_vomap = np.empty(tm * ns, dtype=np.int32)
    seg_rowptr = np.empty((tm + 1) * ns, dtype=np.int32)
    seg_cv = np.empty(2 * nnz, dtype=np.float32)
    tail = np.empty(n_sm + 1, dtype=np.int32)
    nxt = np.empty(n_sm + 1, dtype=np.int32)
    _ffi.call("gcnb_csr2tile", np_ptr(rp), np_ptr(ci), np_ptr(va), m, m, nnz, tm, n_sm,
              np_ptr(seg_vomap), np_ptr(seg_rowptr), np_ptr(seg_cv), np_ptr(tail), np_ptr(nxt),
              ctypes.byref(cnt))
    assert cnt.value == ns
    return dict(seg_rowPtr=seg_rowptr, segNzCV=seg_cv, segVoMap=seg_vomap, n_segs=ns,
                next_seg=nxt, grouped_tailSeg=tail)


def normalize_adj(adj):
    """Â = D^-½ (A + I) D^-½ of a scipy sparse matrix, fp64 arithmetic, returned as CSR fp32 with
    sorted indices — utils.normalize_adj (pygcn/gcnio/util/utils.py:78-90, self-loops added only
    when A[0,0] == 0, exactly as the reference tests it) followed by the fp32 cast of
    sparse_mx_to_torch_sparse_tensor (utils.py:243-250)."""
    import scipy.sparse as sp

    mx = adj.tolil() if not sp.isspmatrix_lil(adj) else adj
    if mx[0, 0] == 0:
        mx = mx + sp.eye(mx.shape[0])
    mx = sp.csr_matrix(mx, dtype=np.float64)
    rowsum = np.asarray(mx.sum(1)).ravel()
    with np.errstate(divide="ignore"):
        r_inv = np.power(rowsum, -0.5)
    r_inv[np.isinf(r_inv)] = 0.0
    d = sp.diags(r_inv)
    out = d.dot(mx).dot(d).tocsr().astype(np.float32)
    out.sort_indices()
    return out

"""TEST INFRASTRUCTURE ONLY — ctypes bindings to the *compiled, unmodified reference*
(oracle/_ref/*.so, built by oracle/Makefile from the sources under /root/reference).

The calls mirror how the reference's own Python binds these libraries
(pygcn/gcn6.py:21-25 LoadLibrary, :313-352 call sites): raw pointers + Python ints, all
functions ``void``.  Only tests/, __graft_entry__.smoke() and bench.py's baseline legs may
import this module.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")


def have(name: str) -> bool:
    return os.path.exists(os.path.join(REF_DIR, name))


_cache = {}


def _lib(name: str):
    if name not in _cache:
        _cache[name] = ctypes.CDLL(os.path.join(REF_DIR, name))
    return _cache[name]


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _csr_copy(rowptr, col, val):
    return (np.ascontiguousarray(rowptr, dtype=np.int32).copy(),
            np.ascontiguousarray(col, dtype=np.int32).copy(),
            np.ascontiguousarray(val, dtype=np.float32).copy())


def _renumber(fn_name, rowptr, col, val, vomp_in=None):
    """dfs / gorder / perm_apply / rabbit (renumber.cu:23,157,233,319): rewrite the CSR in
    place and (except perm_apply, which reads it) fill vomp[new] = old."""
    lib = _lib("ref_graph.so")
    rp, ci, va = _csr_copy(rowptr, col, val)
    m = len(rp) - 1
    vomp = (np.arange(m, dtype=np.int32) if vomp_in is None
            else np.ascontiguousarray(vomp_in, dtype=np.int32).copy())
    getattr(lib, fn_name)(_p(rp), _p(ci), _p(va), _p(vomp), m, m, len(ci))
    return rp, ci, va, vomp


def ref_dfs(rowptr, col, val):
    return _renumber("dfs", rowptr, col, val)


def ref_gorder(rowptr, col, val):
    return _renumber("gorder", rowptr, col, val)


def ref_rabbit(rowptr, col, val):
    return _renumber("rabbit", rowptr, col, val)


def ref_perm_apply(rowptr, col, val, vomp):
    return _renumber("perm_apply", rowptr, col, val, vomp)[:3]


def _order(fn_name, rowptr, col, *extra):
    lib = _lib("ref_graph.so")
    rp = np.ascontiguousarray(rowptr, dtype=np.int32)
    ci = np.ascontiguousarray(col, dtype=np.int32)
    m = len(rp) - 1
    out = np.empty(m, dtype=np.int64)
    getattr(lib, fn_name)(_p(rp), _p(ci), m, len(ci), *extra, _p(out))
    return out


def ref_order_deg(rowptr, col, desc=True):
    """order_deg(Edgelist&, desc) via oracle/ref_shim.cpp; rank[old] = new."""
    return _order("ref_order_deg", rowptr, col, int(bool(desc)))


def ref_order_rcm(rowptr, col):
    return _order("ref_order_rcm", rowptr, col)


def ref_order_gorder(rowptr, col, window=3):
    return _order("ref_order_gorder", rowptr, col, int(window))


def ref_csr2seg(rowptr, col, val, tm=8, corrected=True):
    """Run csr2seg_Cmajor (tile.cu:11-103) panel by panel, the way csr2tile (tile.cu:109-113)
    does, without csr2tile's device query (which needs a GPU, tile.cu:118-122).
    ``corrected`` selects tile_fixed.so (one-token correction at tile.cu:71)."""
    lib = _lib("tile_fixed.so" if corrected else "tile.so")
    rp, ci, va = _csr_copy(rowptr, col, val)
    m = len(rp) - 1
    nnz = len(ci)
    cap = max(nnz, 16) * 9 + 64  # pygcn/gcn6.py:334-336 guesses nnz; be generous for tiny cases
    seg_rowptr = np.zeros(cap, dtype=np.int32)
    seg_cv = np.zeros(2 * nnz + 16, dtype=np.float32)
    seg_vomap = np.zeros(cap, dtype=np.int32)
    vomp = np.arange(m, dtype=np.int32)
    n_segs = np.zeros(1, dtype=np.int32)
    for ridx in range((m + tm - 1) // tm):
        lib.csr2seg_Cmajor(ridx, _p(rp), _p(ci), _p(va), m, m, nnz, _p(vomp), _p(seg_vomap),
                           _p(seg_rowptr), _p(seg_cv), tm, _p(n_segs))
    ns = int(n_segs[0])
    return dict(seg_rowPtr=seg_rowptr[: (tm + 1) * ns].copy(), segNzCV=seg_cv[: 2 * nnz].copy(),
                segVoMap=seg_vomap[: tm * ns].copy(), n_segs=ns)


def ref_csr2tile_gpu(rowptr, col, val, tm=8, corrected=True):
    """Call csr2tile itself (tile.cu:104-169).  Needs a CUDA device for
    cudaGetDeviceProperties (tile.cu:118-122): GPU box only."""
    lib = _lib("tile_fixed.so" if corrected else "tile.so")
    rp, ci, va = _csr_copy(rowptr, col, val)
    m = len(rp) - 1
    nnz = len(ci)
    cap = max(nnz, 16) * 9 + 64
    seg_rowptr = np.zeros(cap, dtype=np.int32)
    seg_cv = np.zeros(2 * nnz + 16, dtype=np.float32)
    seg_vomap = np.zeros(cap, dtype=np.int32)
    tail = np.zeros(1024, dtype=np.int32)
    nxt = np.zeros(1024, dtype=np.int32)
    vomp = np.arange(m, dtype=np.int32)
    n_segs = np.zeros(1, dtype=np.int32)
    lib.csr2tile(_p(rp), _p(ci), _p(va), m, m, nnz, _p(vomp), _p(seg_vomap), _p(seg_rowptr),
                 _p(seg_cv), _p(tail), _p(nxt), tm, _p(n_segs))
    ns = int(n_segs[0])
    return dict(seg_rowPtr=seg_rowptr[: (tm + 1) * ns].copy(), segNzCV=seg_cv[: 2 * nnz].copy(),
                segVoMap=seg_vomap[: tm * ns].copy(), n_segs=ns, next_seg=nxt, grouped_tailSeg=tail)

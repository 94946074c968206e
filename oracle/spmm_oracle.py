"""TEST INFRASTRUCTURE ONLY — ctypes binding of oracle/liboracle.so (the C restatement of the
reference's SpMM arithmetic, oracle/spmm_oracle.c) plus the reference's CPU layer as the reference
itself computes it (torch.spmm on COO tensors, pygcn/gcn1.py:40-58).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  Nothing in gcn_b200/ does.

Parity status: PINNED against outputs of the reference executed in the build container
(tests/golden/make_golden.py → tests/golden/*.npz; checked by tests/test_oracle.py).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")


def build():
    subprocess.run(["make", "-C", _HERE, "port"], check=True, stdout=subprocess.DEVNULL)


def _lib():
    if not os.path.exists(_LIB):
        build()
    return ctypes.CDLL(_LIB)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _csr(rowptr, col, val):
    return (np.ascontiguousarray(rowptr, dtype=np.int32), np.ascontiguousarray(col, dtype=np.int32),
            np.ascontiguousarray(val, dtype=np.float32))


def spmm_f32(rowptr, col, val, n, B):
    """C = A·B, fp32 serial accumulation per output element in ascending column order
    (flexspmm.cu:72-81; torch.spmm semantics, pygcn/gcn1.py:53)."""
    rp, ci, va = _csr(rowptr, col, val)
    B = np.ascontiguousarray(B, dtype=np.float32)
    m, k = len(rp) - 1, B.shape[1]
    C = np.empty((m, k), dtype=np.float32)
    _lib().oracle_spmm_csr_f32(_p(rp), _p(ci), _p(va), m, int(n), k, _p(B), _p(C))
    return C


def spmm_f64(rowptr, col, val, n, B):
    """Same product with fp64 accumulation (the value both fp32 paths are measured against)."""
    rp, ci, va = _csr(rowptr, col, val)
    B = np.ascontiguousarray(B, dtype=np.float32)
    m, k = len(rp) - 1, B.shape[1]
    C = np.empty((m, k), dtype=np.float64)
    _lib().oracle_spmm_csr_f64acc(_p(rp), _p(ci), _p(va), m, int(n), k, _p(B), _p(C))
    return C


def spmm_t_f64(rowptr, col, val, n, B):
    """C[n×k] = Aᵀ·B, fp64 accumulation: the true backward of torch.spmm."""
    rp, ci, va = _csr(rowptr, col, val)
    B = np.ascontiguousarray(B, dtype=np.float32)
    m, k = len(rp) - 1, B.shape[1]
    C = np.empty((int(n), k), dtype=np.float64)
    _lib().oracle_spmm_csr_t_f64acc(_p(rp), _p(ci), _p(va), m, int(n), k, _p(B), _p(C))
    return C


def flexspmm_tiles(tiles, m, n, B, tm=8):
    """The reference kernel's arithmetic over its own tile format (flexspmm.cu:40-93)."""
    B = np.ascontiguousarray(B, dtype=np.float32)
    k = B.shape[1]
    C = np.zeros((m, k), dtype=np.float32)
    srp = np.ascontiguousarray(tiles["seg_rowPtr"], dtype=np.int32)
    scv = np.ascontiguousarray(tiles["segNzCV"], dtype=np.float32)
    svm = np.ascontiguousarray(tiles["segVoMap"], dtype=np.int32)
    _lib().oracle_flexspmm_tiles(_p(srp), _p(scv), _p(svm), int(m), int(n), k, int(tiles["n_segs"]),
                                 tm, _p(B), _p(C))
    return C


def permute_rows(X, vomp):
    """X_out[r] = X[vomp[r]] (permutate.cu:3-59, labels untouched)."""
    X = np.ascontiguousarray(X, dtype=np.float32)
    vo = np.ascontiguousarray(vomp, dtype=np.int32)
    out = np.empty_like(X)
    _lib().oracle_permute_rows(_p(X), _p(vo), X.shape[0], X.shape[1], _p(out))
    return out


# ---- the reference's CPU layer, as the reference computes it ---------------------------------------
def torch_coo(rowptr, col, val, shape):
    """COO fp32 tensor with int64 indices — sparse_mx_to_torch_sparse_tensor,
    pygcn/gcnio/util/utils.py:243-250."""
    import warnings

    import torch
    warnings.filterwarnings("ignore", message="Sparse invariant checks")
    rp = np.asarray(rowptr, dtype=np.int64)
    rows = np.repeat(np.arange(len(rp) - 1, dtype=np.int64), np.diff(rp))
    idx = torch.from_numpy(np.vstack((rows, np.asarray(col, dtype=np.int64))))
    return torch.sparse_coo_tensor(idx, torch.from_numpy(np.asarray(val, dtype=np.float32)),
                                   tuple(shape)).coalesce()


def torch_spmm_cpu(adj_coo, B):
    """output = torch.spmm(adj, support) on the host cores (pygcn/gcn1.py:53, gcn5.py:90)."""
    import torch
    return torch.spmm(adj_coo, B)


def gcn_layer_cpu(adj_coo, x, weight, bias=None):
    """GraphConvolution.forward of the reference on CPU (pygcn/gcn1.py:40-58):
    support = mm/spmm(input, W); output = spmm(adj, support); + bias."""
    import torch
    support = torch.spmm(x, weight) if x.is_sparse else torch.mm(x, weight)
    out = torch.spmm(adj_coo, support)
    return out + bias if bias is not None else out


def gcn_layer2_cpu(adj_coo, x, weight, bias=None):
    """GraphConvolution2.forward of the reference on CPU (pygcn/gcn5.py:134-139; flex form gcn6.py:184-194):
    support = spmm(adj, input); output = mm(support, W); + bias."""
    import torch
    out = torch.mm(torch.spmm(adj_coo, x), weight)
    return out + bias if bias is not None else out


def gcn_epoch_cpu(adj_coo, x, labels, idx_train, params, optimizer, dropout=0.5):
    """One training epoch of the reference's 2-layer GCN on the host cores, restated from
    pygcn/gcn1.py:62-118 (model: gc1 = A·(X·W1)+b1, ReLU, dropout, gc2 = A·(H·W2)+b2, log_softmax) and
    :176-203 (_train_without_val: zero_grad, forward, nll_loss on idx_train, backward, Adam step).
    `params` = (W1, b1, W2, b2) leaf tensors the optimizer owns.  Returns the loss tensor."""
    import torch
    import torch.nn.functional as F
    W1, b1, W2, b2 = params
    optimizer.zero_grad()
    h = F.relu(gcn_layer_cpu(adj_coo, x, W1, b1))
    h = F.dropout(h, dropout, training=True)
    out = F.log_softmax(gcn_layer_cpu(adj_coo, h, W2, b2), dim=1)
    loss = F.nll_loss(out[idx_train], labels[idx_train])
    loss.backward()
    optimizer.step()
    return loss

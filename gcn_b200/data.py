"""On-disk dataset formats the reference reads — host-side, no GPU involved.

Two formats feed the reference's GCN:

* GraphSAINT directories (``adj_full.npz, adj_train.npz, feats.npy, class_map.json, role.json``;
  Reddit, Flickr, Yelp, Amazon, PPI, ogbn-* conversions), read by ``load_data`` +
  ``process_graph_data`` of ``profiling_gcn.py:22-72``;
* DeepRobust / nettack ``<name>.npz`` files (cora, citeseer, cora_ml, polblogs), read by
  ``Dataset.load_npz`` / ``get_adj`` of ``pygcn/gcnio/data/dataio.py:105-150``.

Same files in, same arrays out (``tests/test_data.py`` checks both against fixtures produced by
executing the reference's own functions).  There is no network here and none is attempted: the
reference's download branch (dataio.py:52-66) is replaced by a FileNotFoundError naming the file.
Differences on purpose: paths are explicit (the reference formats ``./{prefix}/...`` relative to the
CWD), and labels always come back as a flat int64 vector (the reference leaves an ``[n, 1]`` float
array for dataset names outside its hard-coded list, profiling_gcn.py:68-69).
"""
from __future__ import annotations

import json
import os

import numpy as np
import scipy.sparse as sp


def _need(path: str) -> str:
    if not os.path.exists(path):
        raise FileNotFoundError(f"gcn_b200.data: {path} not found (datasets are not downloaded; "
                                "place the file there)")
    return path


def load_graphsaint(prefix: str, standardize: bool = True):
    """GraphSAINT directory -> (adj_full CSR, adj_train CSR, feats float64 [n×f], labels int64 [n],
    role dict with 'tr'/'va'/'te' index lists).

    Features are standardised with mean / variance of the TRAINING vertices only — the rows of
    ``adj_train`` that have at least one non-zero — exactly as profiling_gcn.py:31-36 does with
    sklearn's StandardScaler (zero-variance columns are left unscaled).  Labels follow
    process_graph_data (profiling_gcn.py:55-67): a multi-hot / one-hot list per vertex becomes its
    argmax, taken in FILE order of the map (the reference's ``p`` counter), a scalar class id is stored
    at its vertex id."""
    adj_full = sp.load_npz(_need(os.path.join(prefix, "adj_full.npz")))
    adj_train = sp.load_npz(_need(os.path.join(prefix, "adj_train.npz")))
    with open(_need(os.path.join(prefix, "role.json"))) as f:
        role = json.load(f)
    feats = np.load(_need(os.path.join(prefix, "feats.npy")))
    with open(_need(os.path.join(prefix, "class_map.json"))) as f:
        class_map = {int(k): v for k, v in json.load(f).items()}
    if len(class_map) != feats.shape[0]:
        raise ValueError(f"class_map has {len(class_map)} entries, feats has {feats.shape[0]} rows")
    if standardize:
        from sklearn.preprocessing import StandardScaler   # the reference's scaler: same arithmetic
        train_nodes = np.array(list(set(adj_train.nonzero()[0])))
        scaler = StandardScaler()
        scaler.fit(feats[train_nodes])
        feats = scaler.transform(feats)
    n = adj_full.shape[0]
    labels = np.zeros(n, dtype=np.int64)
    first = next(iter(class_map.values()))
    if isinstance(first, list):
        for p, v in enumerate(class_map.values()):
            labels[p] = int(np.argmax(v))
    else:
        for k, v in class_map.items():
            labels[k] = int(v)
    return adj_full.tocsr(), adj_train.tocsr(), feats, labels, role


def load_deeprobust_npz(path: str, require_lcc: bool = True):
    """nettack-style ``.npz`` -> (adj CSR float32 symmetric 0/1 without diagonal, features CSR float32,
    labels) as Dataset.get_adj returns them (dataio.py:105-126): symmetrise, clip to 1, optionally
    keep the largest connected component (setting='nettack'), drop the diagonal."""
    with np.load(_need(path)) as loader:
        adj = sp.csr_matrix((loader["adj_data"], loader["adj_indices"], loader["adj_indptr"]),
                            shape=loader["adj_shape"])
        if "attr_data" in loader:
            features = sp.csr_matrix((loader["attr_data"], loader["attr_indices"], loader["attr_indptr"]),
                                     shape=loader["attr_shape"])
        else:
            features = None
        labels = loader.get("labels")
    if features is None:
        features = np.eye(adj.shape[0])
    features = sp.csr_matrix(features, dtype=np.float32)
    adj = adj + adj.T
    adj = adj.tolil()
    adj[adj > 1] = 1
    if require_lcc:
        keep = largest_connected_component(adj)
        adj = adj[keep][:, keep]
        features = features[keep]
        labels = labels[keep]
        if adj.sum(0).A1.min() <= 0:
            raise ValueError("graph contains singleton nodes")
    adj.setdiag(0)
    adj = adj.astype("float32").tocsr()
    adj.eliminate_zeros()
    if np.abs(adj - adj.T).sum() != 0:
        raise ValueError("input graph is not symmetric")
    if adj.nnz and not (adj.max() == 1 and len(np.unique(adj.data)) == 1):
        raise ValueError("graph must be unweighted")
    return adj, features, labels


def largest_connected_component(adj) -> list:
    """Vertex ids (ascending) of the largest connected component (dataio.py:152-160; ties between
    equally large components resolve like the reference's reversed argsort: the highest label)."""
    _, comp = sp.csgraph.connected_components(adj)
    sizes = np.bincount(comp)
    keep = np.argsort(sizes)[::-1][0]
    return [i for i, c in enumerate(comp) if c == keep]


def get_train_val_test(nnodes: int, val_size=0.1, test_size=0.8, stratify=None, seed=None):
    """The nettack / mettack split of the reference's small graphs — 10 % training, 10 % validation, 80 %
    test, stratified by label (pygcn/gcnio/util/utils.py:270-297, called from dataio.py:36-39 with
    seed 15).  Two sklearn ``train_test_split`` calls drawing from numpy's global generator after
    ``np.random.seed(seed)``, as there, so the same seed gives the same three index arrays."""
    from sklearn.model_selection import train_test_split

    if stratify is None:
        raise ValueError("stratify cannot be None")
    if seed is not None:
        np.random.seed(seed)
    idx = np.arange(nnodes)
    train_size = 1 - val_size - test_size
    idx_train_and_val, idx_test = train_test_split(idx, random_state=None, train_size=train_size + val_size,
                                                   test_size=test_size, stratify=stratify)
    stratify = np.asarray(stratify)[idx_train_and_val]
    idx_train, idx_val = train_test_split(idx_train_and_val, random_state=None,
                                          train_size=train_size / (train_size + val_size),
                                          test_size=val_size / (train_size + val_size), stratify=stratify)
    return idx_train, idx_val, idx_test


def split_indices(role: dict):
    """(idx_train, idx_val, idx_test) int64 arrays from a GraphSAINT role map (profiling_gcn.py:131-133)."""
    return (np.asarray(role["tr"], dtype=np.int64), np.asarray(role["va"], dtype=np.int64),
            np.asarray(role["te"], dtype=np.int64))

"""Command-line driver of the GCN path — the role of the reference's ``profiling_gcn.py:74-172``.

    python -m gcn_b200.profiling -g reddit -k 128 [-i 100] [--root ./dataset] [--reorder rabbit]
                                 [--association auto] [--cuda-graph] [--test]

Same switches as the reference (``-g/--graph``, ``-k/--hidden``, ``-i/--train-iters`` default 100,
``profiling_gcn.py:85-90``), same seed (15, ``:75-79``), same flow: load the graph, build
``GCN(nfeat, nhid=k, nclass, dataset, device)`` (``:155``), ``fit(features, adj, labels, idx_train,
train_iters, verbose=True, name=dataset)`` (``:164``).  Graphs: a GraphSAINT directory
``<root>/<graph>/`` (the reference's "large graphs": flickr, ppi, amazon, reddit, yelp, ``:82``) or a
nettack ``<root>/<graph>.npz`` (its "small graphs").  Nothing is downloaded.  The reference accepts
only the names of its two hard-coded sets and exits 1 on anything else; here any name that resolves to
one of the two on-disk formats is accepted, and a name that resolves to neither exits 1 with the two
paths that were tried.

Additions, all off by default: ``--reorder`` (the reference selects its reordering with module
constants, ``gcn6.py:27-30``), ``--association``, ``--cuda-graph``, ``--test`` (the reference's
``TEST`` branch, ``:168-172``, evaluated on the model just trained instead of a saved checkpoint).
Prints one summary line at the end: vertices, non-zeros, widths, and the mean epoch time measured
with CUDA events around the whole ``fit`` loop.  CUDA only — there is no CPU path.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

SEED = 15   # profiling_gcn.py:75


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser("python -m gcn_b200.profiling", description="Graph to be processed ...")
    ap.add_argument("-g", "--graph", required=True, help="dataset name under --root, or a path")
    ap.add_argument("-k", "--hidden", type=int, required=True)
    ap.add_argument("-i", "--train-iters", dest="train_iters", type=int, default=100)
    ap.add_argument("--root", default="./dataset", help="directory holding <graph>/ or <graph>.npz")
    ap.add_argument("--reorder", default=None, choices=["rabbit", "gorder", "rcm", "deg", "dfs"])
    ap.add_argument("--association", default="reference", choices=["reference", "auto"])
    ap.add_argument("--cuda-graph", action="store_true")
    ap.add_argument("--test", action="store_true", help="evaluate on the test split after training")
    ap.add_argument("--device", default="cuda:0")
    return ap


def resolve(graph: str, root: str):
    """-> ('graphsaint', directory) | ('npz', file) for a dataset name or explicit path; None if neither."""
    cands = [graph] if os.path.sep in graph or os.path.exists(graph) else []
    cands.append(os.path.join(root, graph))
    tried = []
    for c in cands:
        if os.path.isdir(c) and os.path.exists(os.path.join(c, "adj_full.npz")):
            return ("graphsaint", c), tried
        tried.append(os.path.join(c, "adj_full.npz"))
        f = c if c.endswith(".npz") else c + ".npz"
        if os.path.isfile(f):
            return ("npz", f), tried
        tried.append(f)
    return None, tried


def load(kind: str, path: str):
    """-> (adj scipy CSR, features scipy CSR, labels int64, idx_train, idx_val, idx_test)."""
    import scipy.sparse as sp

    from . import data

    if kind == "graphsaint":
        adj, _adj_train, feats, labels, role = data.load_graphsaint(path)
        idx_train, idx_val, idx_test = data.split_indices(role)
        return adj, sp.csr_matrix(feats), labels, idx_train, idx_val, idx_test     # profiling_gcn.py:129
    adj, feats, labels = data.load_deeprobust_npz(path)
    if labels is None:
        raise ValueError(f"{path} holds no labels")
    labels = np.asarray(labels).astype(np.int64)
    # Dataset(setting='nettack', seed=15): 10 % / 10 % / 80 % stratified split (dataio.py:36-39)
    idx_train, idx_val, idx_test = data.get_train_val_test(adj.shape[0], stratify=labels, seed=SEED)
    return adj, feats, labels, idx_train, idx_val, idx_test


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    found, tried = resolve(args.graph, args.root)
    if found is None:
        print("Except input graph ... (looked for " + ", ".join(tried) + ")")     # profiling_gcn.py:99-100
        return 1
    import torch

    if not torch.cuda.is_available():
        raise RuntimeError("gcn_b200.profiling: no CUDA device (there is no CPU path)")
    from .model import GCN

    np.random.seed(SEED)
    torch.manual_seed(SEED)
    torch.cuda.manual_seed(SEED)
    dataset = os.path.basename(os.path.normpath(found[1])).removesuffix(".npz")
    print(dataset)
    adj, features, labels, idx_train, idx_val, idx_test = load(*found)
    nclass = int(labels.max()) + 1
    print("labels = ", nclass)
    device = torch.device(args.device)
    model = GCN(nfeat=features.shape[1], nhid=args.hidden, nclass=nclass, dataset=dataset, device=device,
                association=args.association).to(device)
    # the reference hands the features over as a scipy CSR matrix and densifies them on the device
    # (gcn6.py:358); so does this: dense fp32 rows
    x = torch.from_numpy(np.asarray(features.todense(), dtype=np.float32))
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(device)
    t0.record()
    model.fit(x, adj, labels, idx_train, train_iters=args.train_iters, verbose=True, name=dataset,
              reorder=args.reorder, cuda_graph=args.cuda_graph)
    t1.record()
    torch.cuda.synchronize(device)
    ms = t0.elapsed_time(t1)
    print(f"{dataset}: n={adj.shape[0]} nnz(adj)={adj.nnz} nfeat={features.shape[1]} nhid={args.hidden} "
          f"nclass={nclass} train_iters={args.train_iters} fit={ms:.1f} ms "
          f"({ms / max(1, args.train_iters):.3f} ms/epoch incl. preprocessing)")
    if args.test:
        model.test(idx_test)
    return 0


if __name__ == "__main__":
    sys.exit(main())

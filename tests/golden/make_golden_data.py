"""Fixtures for tests/test_data.py, produced by EXECUTING the reference's own loader functions
(profiling_gcn.py:22-72 load_data / process_graph_data, extracted by name with `ast` because the module
runs a training job at import; pygcn/gcnio/data/dataio.py Dataset.get_adj / load_npz, called unbound)
on small seeded files that `write_inputs` below regenerates identically in the test.
Run in the build container (needs /root/reference):  python tests/golden/make_golden_data.py"""
import ast
import json
import os
import sys
import tempfile

import numpy as np
import scipy.sparse as sp

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def write_inputs(root, labels_as_lists):
    """A GraphSAINT-style directory `root/saint` and a nettack-style `root/toy.npz`."""
    rng = np.random.default_rng(11)
    n, f, c = 60, 7, 5
    a = sp.random(n, n, density=0.08, format="csr", random_state=np.random.RandomState(11), dtype=np.float64)
    a.data[:] = 1.0
    adj_full = ((a + a.T) > 0).astype(np.float64).tocsr()
    train = np.sort(rng.choice(n, 30, replace=False))
    mask = np.zeros(n, bool); mask[train] = True
    adj_train = adj_full.multiply(mask[:, None]).multiply(mask[None, :]).tocsr()
    feats = rng.standard_normal((n, f)) * np.arange(1, f + 1) + 3.0
    feats[:, 2] = 4.25                                  # zero variance column
    y = rng.integers(0, c, n)
    if labels_as_lists:
        class_map = {str(i): np.eye(c, dtype=int)[y[i]].tolist() for i in rng.permutation(n)}   # shuffled file order
    else:
        class_map = {str(i): int(y[i]) for i in range(n)}
    rest = np.setdiff1d(np.arange(n), train)
    role = {"tr": train.tolist(), "va": rest[:10].tolist(), "te": rest[10:].tolist()}
    d = os.path.join(root, "saint")
    os.makedirs(d, exist_ok=True)
    sp.save_npz(os.path.join(d, "adj_full.npz"), adj_full)
    sp.save_npz(os.path.join(d, "adj_train.npz"), adj_train)
    np.save(os.path.join(d, "feats.npy"), feats)
    json.dump(class_map, open(os.path.join(d, "class_map.json"), "w"))
    json.dump(role, open(os.path.join(d, "role.json"), "w"))
    # nettack-style npz: directed, weighted duplicates, self loops, two components + an isolated pair
    m = 40
    b = sp.random(m, m, density=0.06, format="csr", random_state=np.random.RandomState(5), dtype=np.float64)
    b.data[:] = rng.integers(1, 3, b.nnz)
    b = b.tolil(); b[30:, :30] = 0; b[:30, 30:] = 0; b[3, 3] = 1; b = b.tocsr()
    attr = sp.random(m, 9, density=0.3, format="csr", random_state=np.random.RandomState(6), dtype=np.float32)
    np.savez(os.path.join(root, "toy.npz"), adj_data=b.data, adj_indices=b.indices, adj_indptr=b.indptr,
             adj_shape=b.shape, attr_data=attr.data, attr_indices=attr.indices, attr_indptr=attr.indptr,
             attr_shape=attr.shape, labels=rng.integers(0, 3, m))
    return d, os.path.join(root, "toy.npz")


def reference_functions():
    src = open(os.path.join(REF, "profiling_gcn.py")).read()
    tree = ast.parse(src)
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("load_data", "process_graph_data")]
    ns = {}
    exec("import json, scipy.sparse, numpy as np\nfrom numpy import argmax\nimport scipy\n"
         "from sklearn.preprocessing import StandardScaler\n", ns)
    exec(compile(ast.Module(body=keep, type_ignores=[]), "profiling_gcn.py", "exec"), ns)
    return ns["load_data"], ns["process_graph_data"]


def main():
    sys.path.insert(0, REF)
    from pygcn.gcnio.data import dataio
    load_data, process_graph_data = reference_functions()
    out = {}
    for as_lists in (False, True):
        with tempfile.TemporaryDirectory() as root:
            d, npz = write_inputs(root, as_lists)
            cwd = os.getcwd()
            os.chdir(root)                      # the reference formats './{prefix}/...'
            try:
                res = process_graph_data(*load_data("saint"), "reddit")
            finally:
                os.chdir(cwd)
            adj_full, adj_train, feats, class_arr, role = res
            tag = "lists" if as_lists else "ints"
            out[f"saint_{tag}_feats"] = feats
            out[f"saint_{tag}_labels"] = np.asarray(class_arr)
            out[f"saint_{tag}_adj_indptr"], out[f"saint_{tag}_adj_indices"] = adj_full.tocsr().indptr, adj_full.tocsr().indices
            if not as_lists:
                for lcc in (True, False):
                    ds = dataio.Dataset.__new__(dataio.Dataset)
                    ds.data_filename, ds.require_lcc = npz, lcc
                    adj, features, labels = ds.get_adj()
                    k = f"npz_lcc{int(lcc)}"
                    out[k + "_indptr"], out[k + "_indices"], out[k + "_data"] = adj.indptr, adj.indices, adj.data
                    f = features.tocsr(); f.sort_indices()
                    out[k + "_f_indptr"], out[k + "_f_indices"], out[k + "_f_data"] = f.indptr, f.indices, f.data
                    out[k + "_labels"] = np.asarray(labels)
    # nettack split of the small graphs (utils.py:270-297), seed 15 as profiling_gcn.py / dataio.py use
    from pygcn.gcnio.util.utils import get_train_val_test
    lab = np.random.default_rng(4).integers(0, 4, 200)
    out["split_labels"] = lab
    out["split_train"], out["split_val"], out["split_test"] = get_train_val_test(200, stratify=lab, seed=15)
    np.savez_compressed(os.path.join(HERE, "loaders.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()

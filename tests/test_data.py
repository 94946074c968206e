"""Dataset loaders against fixtures produced by executing the reference's own
functions on the same seeded files (tests/golden/make_golden_data.py): GraphSAINT directories
(profiling_gcn.py:22-72) and nettack-style .npz (pygcn/gcnio/data/dataio.py:105-160).  CPU only."""
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN
from gcn_b200 import data

sys.path.insert(0, GOLDEN)
from make_golden_data import write_inputs  # noqa: E402  (input generator only; no reference code)


@pytest.fixture(scope="module")
def golden():
    return dict(np.load(os.path.join(GOLDEN, "loaders.npz")))


@pytest.mark.parametrize("as_lists", [False, True])
def test_graphsaint_directory_matches_reference(tmp_path, golden, as_lists):
    d, _ = write_inputs(str(tmp_path), as_lists)
    adj_full, adj_train, feats, labels, role = data.load_graphsaint(d)
    tag = "lists" if as_lists else "ints"
    assert np.array_equal(feats, golden[f"saint_{tag}_feats"])           # same scaler arithmetic: bit-exact
    assert labels.dtype == np.int64 and np.array_equal(labels, golden[f"saint_{tag}_labels"])
    assert np.array_equal(adj_full.indptr, golden[f"saint_{tag}_adj_indptr"])
    assert np.array_equal(adj_full.indices, golden[f"saint_{tag}_adj_indices"])
    assert abs(feats[role["tr"]].mean()) < 1 and np.allclose(feats[role["tr"]][:, 0].std(), 1.0, atol=0.2)
    assert np.all(feats[:, 2] == 0)                                       # zero-variance column: centred, unscaled
    tr, va, te = data.split_indices(role)
    assert len(tr) + len(va) + len(te) == adj_full.shape[0] and adj_train.nnz <= adj_full.nnz
    raw = data.load_graphsaint(d, standardize=False)[2]
    assert np.array_equal(raw, np.load(os.path.join(d, "feats.npy")))


@pytest.mark.parametrize("lcc", [True, False])
def test_deeprobust_npz_matches_reference(tmp_path, golden, lcc):
    _, npz = write_inputs(str(tmp_path), False)
    adj, features, labels = data.load_deeprobust_npz(npz, require_lcc=lcc)
    k = f"npz_lcc{int(lcc)}"
    assert adj.dtype == np.float32
    assert np.array_equal(adj.indptr, golden[k + "_indptr"]) and np.array_equal(adj.indices, golden[k + "_indices"])
    assert np.array_equal(adj.data, golden[k + "_data"])
    f = features.tocsr()
    f.sort_indices()
    assert np.array_equal(f.indptr, golden[k + "_f_indptr"]) and np.array_equal(f.indices, golden[k + "_f_indices"])
    assert np.array_equal(f.data, golden[k + "_f_data"])
    assert np.array_equal(np.asarray(labels), golden[k + "_labels"])
    assert (adj != adj.T).nnz == 0 and adj.diagonal().sum() == 0


def test_missing_files_are_reported_not_downloaded(tmp_path):
    with pytest.raises(FileNotFoundError, match="not downloaded"):
        data.load_graphsaint(str(tmp_path / "nowhere"))
    with pytest.raises(FileNotFoundError):
        data.load_deeprobust_npz(str(tmp_path / "cora.npz"))


def test_loaded_graph_feeds_the_host_pipeline(tmp_path):
    """Loader -> normalize_adj -> reorder: the arrays have the dtypes/layout the C ABI takes."""
    from gcn_b200 import graph
    d, _ = write_inputs(str(tmp_path), False)
    adj_full, _, feats, labels, _ = data.load_graphsaint(d)
    A = graph.normalize_adj(adj_full)
    rp, ci, va, vomp = graph.renumber_rabbit(A.indptr, A.indices, A.data)
    assert sorted(vomp.tolist()) == list(range(A.shape[0])) and rp[-1] == A.nnz
    assert feats[vomp].shape == feats.shape and labels[vomp].shape == labels.shape


def test_nettack_split_matches_reference(golden):
    """10 % / 10 % / 80 % stratified split of the small graphs (utils.py:270-297) with the reference's
    seed: same three index arrays as the reference's function returned for the same labels."""
    lab = golden["split_labels"]
    tr, va, te = data.get_train_val_test(len(lab), stratify=lab, seed=15)
    assert np.array_equal(tr, golden["split_train"])
    assert np.array_equal(va, golden["split_val"])
    assert np.array_equal(te, golden["split_test"])
    assert len(np.unique(np.concatenate([tr, va, te]))) == len(lab)
    with pytest.raises(ValueError):
        data.get_train_val_test(10, stratify=None)

"""Command-line driver (gcn_b200/profiling.py; the role of the reference's profiling_gcn.py:74-172):
argument surface, dataset resolution and loading on CPU; one short training run on the GPU.
Named to run last: it drives everything the other files test piece by piece."""
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN
from gcn_b200 import profiling

sys.path.insert(0, GOLDEN)
from make_golden_data import write_inputs  # noqa: E402  (input generator only; no reference code)


def test_reference_switches_and_defaults():
    a = profiling.build_parser().parse_args(["-g", "reddit", "-k", "128"])     # profiling_gcn.py:85-90
    assert (a.graph, a.hidden, a.train_iters) == ("reddit", 128, 100)
    a = profiling.build_parser().parse_args(["--graph", "yelp", "--hidden", "64", "--train-iters", "7"])
    assert (a.graph, a.hidden, a.train_iters, a.reorder, a.cuda_graph) == ("yelp", 64, 7, None, False)
    assert profiling.SEED == 15                                                # profiling_gcn.py:75


def test_resolves_both_formats_and_rejects_unknown(tmp_path, capsys):
    d, npz = write_inputs(str(tmp_path), False)
    assert profiling.resolve("saint", str(tmp_path))[0] == ("graphsaint", d)
    assert profiling.resolve("toy", str(tmp_path))[0] == ("npz", npz)
    assert profiling.resolve(d, "/nowhere")[0] == ("graphsaint", d)            # explicit paths
    assert profiling.resolve(npz, "/nowhere")[0] == ("npz", npz)
    found, tried = profiling.resolve("pubmed", str(tmp_path))
    assert found is None and len(tried) == 2
    assert profiling.main(["-g", "pubmed", "-k", "16", "--root", str(tmp_path)]) == 1   # exit(1), :99-100
    assert "Except input graph" in capsys.readouterr().out


def test_load_graphsaint_directory(tmp_path):
    d, _ = write_inputs(str(tmp_path), False)
    adj, feats, labels, tr, va, te = profiling.load("graphsaint", d)
    n = adj.shape[0]
    assert feats.shape == (n, 7) and labels.shape == (n,) and labels.dtype == np.int64
    assert len(tr) + len(va) + len(te) == n and len(np.intersect1d(tr, te)) == 0


def test_no_cpu_path(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a box without a GPU")
    write_inputs(str(tmp_path), False)
    with pytest.raises(RuntimeError, match="no CUDA device"):
        profiling.main(["-g", "saint", "-k", "16", "--root", str(tmp_path)])


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [[], ["--reorder", "rabbit", "--association", "auto", "--test"]])
def test_driver_trains_on_a_graphsaint_directory(tmp_path, capsys, extra):
    write_inputs(str(tmp_path), False)
    assert profiling.main(["-g", "saint", "-k", "16", "-i", "12", "--root", str(tmp_path)] + extra) == 0
    out = capsys.readouterr().out
    assert out.startswith("saint\n") and "labels =  5" in out and "ms/epoch" in out
    if "--test" in extra:
        assert "Test set results" in out

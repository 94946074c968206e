import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
GRAPH_CASES = ["er_small", "er_ragged", "powerlaw", "banded", "star"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, f"{name}.npz")))


@pytest.fixture(params=GRAPH_CASES)
def golden_graph(request):
    g = load_golden(f"graph_{request.param}")
    g["name"] = request.param
    return g


@pytest.fixture(scope="session")
def golden_gcn():
    return load_golden("gcn_cora")


def random_normalized_graph(n, e, seed, kind="er"):
    """Small seeded Â (scipy CSR fp32) through gcn_b200.graph.normalize_adj."""
    import scipy.sparse as sp
    from gcn_b200 import graph
    rng = np.random.default_rng(seed)
    if kind == "er":
        r, c = rng.integers(0, n, e), rng.integers(0, n, e)
    else:  # skewed
        r = (n * rng.random(e) ** 3).astype(np.int64)
        c = rng.integers(0, n, e)
    a = sp.coo_matrix((np.ones(e), (r, c)), shape=(n, n)).tocsr()
    a = ((a + a.T) > 0).astype(np.float64)
    a.setdiag(0)
    a.eliminate_zeros()
    return graph.normalize_adj(a)


def rel_err(x, ref):
    """max |x - ref| / max |ref| (the 1e-5 relative bound of BASELINE.json's north_star)."""
    x = np.asarray(x, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    denom = max(float(np.abs(ref).max()), 1e-30) if ref.size else 1.0
    return float(np.abs(x - ref).max() / denom) if ref.size else 0.0

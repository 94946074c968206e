"""Size-independent properties of the host preprocessing (hypothesis-generated graphs, CPU): every
reordering returns a permutation, the renumbered CSR is exactly P·Â·Pᵀ with ascending columns and
untouched values, and aggregation commutes with it: (PÂPᵀ)(PX) = P(ÂX)."""
import numpy as np
import scipy.sparse as sp
from hypothesis import given, settings
from hypothesis import strategies as st

from gcn_b200 import graph
from oracle import spmm_oracle as SO

RENUMBER = {"deg": graph.renumber_deg, "rcm": graph.renumber_rcm, "gorder": graph.renumber_gorder,
            "dfs": graph.renumber_dfs, "rabbit": graph.renumber_rabbit}


@st.composite
def normalized_graphs(draw):
    n = draw(st.integers(2, 60))
    e = draw(st.integers(0, 4 * n))
    seed = draw(st.integers(0, 2**31 - 1))
    rng = np.random.default_rng(seed)
    r, c = rng.integers(0, n, e), rng.integers(0, n, e)
    a = sp.coo_matrix((np.ones(e), (r, c)), shape=(n, n)).tocsr()
    a = ((a + a.T) > 0).astype(np.float64)
    a.setdiag(0)
    a.eliminate_zeros()
    return graph.normalize_adj(a), seed      # self loops present: no isolated vertex (Gorder's precondition)


@settings(max_examples=40, deadline=None)
@given(normalized_graphs(), st.sampled_from(sorted(RENUMBER)))
def test_renumbering_is_a_permutation_similarity(gs, name):
    A, seed = gs
    n = A.shape[0]
    rp, ci, va, vomp = RENUMBER[name](A.indptr, A.indices, A.data)
    assert sorted(vomp.tolist()) == list(range(n))
    B = sp.csr_matrix((va, ci, rp), shape=(n, n))
    want = A[vomp][:, vomp].tocsr()
    want.sort_indices()
    assert np.array_equal(rp, want.indptr) and np.array_equal(ci, want.indices)
    assert np.array_equal(va, want.data)                       # values moved, never recomputed
    for i in range(n):
        assert np.all(np.diff(ci[rp[i]:rp[i + 1]]) > 0)        # ascending columns: the plan's precondition
    X = np.random.default_rng(seed).standard_normal((n, 3)).astype(np.float32)
    lhs = SO.spmm_f64(rp, ci, va, n, X[vomp])
    rhs = SO.spmm_f64(A.indptr, A.indices, A.data, n, X)[vomp]
    assert np.allclose(lhs, rhs, rtol=0, atol=1e-12)
    assert B.nnz == A.nnz


@settings(max_examples=25, deadline=None)
@given(normalized_graphs())
def test_orders_are_permutations_and_tiles_cover_every_nonzero(gs):
    A, _ = gs
    n = A.shape[0]
    for rank in (graph.order_deg(A.indptr, A.indices, True), graph.order_deg(A.indptr, A.indices, False),
                 graph.order_rcm(A.indptr, A.indices), graph.order_gorder(A.indptr, A.indices, 3)):
        assert sorted(np.asarray(rank).tolist()) == list(range(n))
    t = graph.csr2tile(A.indptr, A.indices, A.data, n_sm=4)
    ns = int(t["n_segs"])
    srp = np.asarray(t["seg_rowPtr"]).reshape(ns, 9) if ns else np.zeros((0, 9), dtype=np.int32)
    assert ns == 0 or (srp[0, 0] == 0 and srp[-1, -1] == A.nnz)     # cumulative offsets cover all non-zeros
    assert np.all(np.diff(srp, axis=1) >= 0)
    cv = np.asarray(t["segNzCV"])[: 2 * A.nnz].reshape(-1, 2)
    assert np.isclose(cv[:, 1].sum(), A.data.sum(), rtol=1e-5)        # every value appears once

"""Host graph preprocessing of libgcn_b200.so (through the C ABI, gcn_b200.graph) — bit-exact
against the reference's outputs in tests/golden/ and, where built, the compiled reference.  CPU."""
import ctypes

import numpy as np
import pytest

from conftest import random_normalized_graph
from gcn_b200 import _ffi, graph
from oracle import ref_bindings as R


def test_orders_bit_exact(golden_graph):
    g = golden_graph
    rp, ci = g["rowptr"], g["col"]
    assert np.array_equal(graph.order_deg(rp, ci, True), g["rank_deg_desc"])
    assert np.array_equal(graph.order_deg(rp, ci, False), g["rank_deg_asc"])
    assert np.array_equal(graph.order_rcm(rp, ci), g["rank_rcm"])
    assert np.array_equal(graph.order_gorder(rp, ci, 3), g["rank_gorder3"])
    assert np.array_equal(graph.order_gorder(rp, ci, 5), g["rank_gorder5"])


def test_renumber_bit_exact(golden_graph):
    g = golden_graph
    rp, ci, va = g["rowptr"], g["col"], g["val"]
    for nm, fn in (("gorder", graph.renumber_gorder), ("dfs", graph.renumber_dfs),
                   ("rabbit", graph.renumber_rabbit)):
        r2, c2, v2, vomp = fn(rp, ci, va)
        assert np.array_equal(vomp, g[f"{nm}_vomp"]), nm
        assert np.array_equal(r2, g[f"{nm}_rowptr"]) and np.array_equal(c2, g[f"{nm}_col"])
        assert np.array_equal(v2, g[f"{nm}_val"])
    r2, c2, v2 = graph.perm_apply(rp, ci, va, g["perm"])
    assert np.array_equal(r2, g["perm_rowptr"]) and np.array_equal(c2, g["perm_col"])
    assert np.array_equal(v2, g["perm_val"])


def test_rcm_deg_renumber_consistent(golden_graph):
    g = golden_graph
    rp, ci, va = g["rowptr"], g["col"], g["val"]
    for fn, rank in ((graph.renumber_rcm, g["rank_rcm"]),
                     (lambda *a: graph.renumber_deg(*a, desc=True), g["rank_deg_desc"])):
        r2, c2, v2, vomp = fn(rp, ci, va)
        inv = np.empty_like(vomp)
        inv[rank] = np.arange(len(rank))
        assert np.array_equal(vomp, inv)          # vomp[new] = old  (renumber.cu:184-189)
        e = graph.perm_apply(rp, ci, va, vomp)
        assert all(np.array_equal(x, y) for x, y in zip((r2, c2, v2), e))


def test_tiles_bit_exact(golden_graph):
    """n_segs and segVoMap equal the reference as shipped; seg_rowPtr / segNzCV equal the
    reference with the tile.cu:71 one-token correction (cumulative row pointer)."""
    g = golden_graph
    t = graph.csr2tile(g["rowptr"], g["col"], g["val"], n_sm=148)
    assert t["n_segs"] == int(g["tile_n_segs"])
    assert np.array_equal(t["segVoMap"], g["tile_segVoMap"])
    assert np.array_equal(t["seg_rowPtr"], g["tile_seg_rowPtr"])
    assert np.array_equal(t["segNzCV"], g["tile_segNzCV"])
    # buckets (tile.cu:128-161): contiguous, the overflow bucket ends at n_segs; with fewer
    # segments than SMs the reference lets the heads run past n_segs while clamping the tails
    nxt, tail = t["next_seg"], t["grouped_tailSeg"]
    assert nxt[0] == 0 and tail[-1] == t["n_segs"] and np.all(tail <= t["n_segs"])
    live = nxt < t["n_segs"]
    assert np.all(tail[live] > nxt[live])


def test_tile_buckets_match_restatement(golden_graph):
    from oracle import graph_oracle as GO
    g = golden_graph
    n = len(g["rowptr"]) - 1
    for n_sm in (4, 148):
        t = graph.csr2tile(g["rowptr"], g["col"], g["val"], n_sm=n_sm)
        o = GO.csr2tile(g["rowptr"], g["col"], g["val"], n, n_sm)
        assert np.array_equal(t["next_seg"], np.asarray(o["next_seg"])[: n_sm + 1])
        assert np.array_equal(t["grouped_tailSeg"], np.asarray(o["grouped_tailSeg"])[: n_sm + 1])


@pytest.mark.skipif(not R.have("ref_graph.so"), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("seed", range(12))
def test_vs_compiled_reference_random(seed, capfd):
    n = 30 + 97 * seed
    A = random_normalized_graph(n, 4 * n + 50 * seed, 100 + seed, "er" if seed % 3 else "skew")
    rp, ci, va = A.indptr, A.indices, A.data
    for desc in (True, False):
        assert np.array_equal(graph.order_deg(rp, ci, desc), R.ref_order_deg(rp, ci, desc))
    assert np.array_equal(graph.order_rcm(rp, ci), R.ref_order_rcm(rp, ci))
    for w in (1, 3, 7):
        assert np.array_equal(graph.order_gorder(rp, ci, w), R.ref_order_gorder(rp, ci, w))
    for ours, ref in ((graph.renumber_gorder, R.ref_gorder), (graph.renumber_dfs, R.ref_dfs),
                      (graph.renumber_rabbit, R.ref_rabbit)):
        assert all(np.array_equal(x, y) for x, y in zip(ours(rp, ci, va), ref(rp, ci, va)))
    # Rabbit on asymmetric patterns (its degree bookkeeping depends on the direction, renumber.cu:383-399)
    import scipy.sparse as sp
    for tri in (sp.triu, sp.tril):
        T = tri(A).tocsr()
        T.sort_indices()
        assert all(np.array_equal(x, y) for x, y in zip(graph.renumber_rabbit(T.indptr, T.indices, T.data),
                                                        R.ref_rabbit(T.indptr, T.indices, T.data)))
    t, r = graph.csr2tile(rp, ci, va), R.ref_csr2seg(rp, ci, va)
    for key in ("n_segs", "segVoMap", "seg_rowPtr", "segNzCV"):
        assert np.array_equal(t[key], r[key]), key
    capfd.readouterr()


def test_normalize_adj_matches_reference_values(golden_gcn):
    import scipy.sparse as sp
    g = golden_gcn
    n = len(g["a_rowptr"]) - 1
    a = sp.csr_matrix((np.ones(len(g["a_col"])), g["a_col"], g["a_rowptr"]), shape=(n, n))
    ours = graph.normalize_adj(a).tocoo()
    assert np.array_equal(np.vstack((ours.row, ours.col)), g["adj_norm_idx"])
    assert np.array_equal(ours.data, g["adj_norm_val"])   # bit-exact fp32 values


def test_edge_cases_and_errors():
    rp = np.zeros(1, dtype=np.int32)
    ci = np.zeros(0, dtype=np.int32)
    assert len(graph.order_deg(rp, ci)) == 0 and len(graph.order_rcm(rp, ci)) == 0
    # isolated vertex: Gorder is undefined in the reference (unitheap.cu:31-33) -> clean error
    rp = np.array([0, 1, 1, 2], dtype=np.int32)
    ci = np.array([0, 2], dtype=np.int32)
    with pytest.raises(_ffi.GcnB200Error):
        graph.order_gorder(rp, ci)
    # malformed CSR / permutation
    with pytest.raises(_ffi.GcnB200Error):
        graph.order_rcm(np.array([0, 2, 1], dtype=np.int32), np.array([0, 1], dtype=np.int32))
    with pytest.raises(_ffi.GcnB200Error):
        graph.order_deg(np.array([0, 1], dtype=np.int32), np.array([5], dtype=np.int32))
    with pytest.raises(_ffi.GcnB200Error):
        graph.perm_apply(np.array([0, 1, 2], dtype=np.int32), np.array([0, 1], dtype=np.int32),
                         np.ones(2, dtype=np.float32), np.array([0, 0], dtype=np.int32))
    # single vertex with a self loop
    r = graph.renumber_gorder(np.array([0, 1], dtype=np.int32), np.array([0], dtype=np.int32),
                              np.ones(1, dtype=np.float32))
    assert r[3].tolist() == [0]


def test_compat_shim_host_entry_points(golden_graph):
    """The reference's own symbol names and argument orders (renumber.cu:157,233; tile.cu:11-12),
    bound the way pygcn/gcn6.py:313-352 binds them: raw pointers and Python ints."""
    g = golden_graph
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    rp, ci, va = g["rowptr"].copy(), g["col"].copy(), g["val"].copy()
    m, nnz = len(rp) - 1, len(ci)
    vomp = np.zeros(m, dtype=np.int32)
    _ffi.lib.gorder(p(rp), p(ci), p(va), p(vomp), m, m, nnz)
    assert np.array_equal(vomp, g["gorder_vomp"]) and np.array_equal(ci, g["gorder_col"])
    rp, ci, va = g["rowptr"].copy(), g["col"].copy(), g["val"].copy()
    _ffi.lib.dfs(p(rp), p(ci), p(va), p(vomp), m, m, nnz)
    assert np.array_equal(vomp, g["dfs_vomp"]) and np.array_equal(rp, g["dfs_rowptr"])
    rp, ci, va = g["rowptr"].copy(), g["col"].copy(), g["val"].copy()
    _ffi.lib.rabbit(p(rp), p(ci), p(va), p(vomp), m, m, nnz)     # renumber.cu:319, the reference's default order
    assert np.array_equal(vomp, g["rabbit_vomp"]) and np.array_equal(ci, g["rabbit_col"])
    # csr2seg_Cmajor panel by panel, like csr2tile's loop (tile.cu:109-113)
    rp, ci, va = g["rowptr"], g["col"], g["val"]
    ns = int(g["tile_n_segs"])
    svm = np.zeros(8 * ns, dtype=np.int32)
    srp = np.zeros(9 * ns, dtype=np.int32)
    scv = np.zeros(2 * nnz, dtype=np.float32)
    cnt = np.zeros(1, dtype=np.int32)
    ident = np.arange(m, dtype=np.int32)
    for ridx in range((m + 7) // 8):
        _ffi.lib.csr2seg_Cmajor(ridx, p(rp), p(ci), p(va), m, m, nnz, p(ident), p(svm), p(srp), p(scv), 8, p(cnt))
    assert cnt[0] == ns and np.array_equal(svm, g["tile_segVoMap"])
    assert np.array_equal(srp, g["tile_seg_rowPtr"]) and np.array_equal(scv, g["tile_segNzCV"])


def test_threaded_row_loops_change_nothing(monkeypatch):
    """perm_apply and the relabelled adjacency lists sort their rows on several host threads once the
    graph is large enough (graph.h: parallel_rows); the result must be the single-threaded one, and
    (where oracle/_ref is built) the reference's."""
    A = random_normalized_graph(4000, 300_000, 77, "skew")      # ~600k non-zeros: 8+ chunks of 64k
    rp, ci, va = A.indptr, A.indices, A.data
    assert rp[-1] > 8 * (1 << 16)
    vomp = np.random.default_rng(5).permutation(A.shape[0]).astype(np.int32)
    got = {}
    for threads in ("1", "5", "16"):
        monkeypatch.setenv("GCNB_HOST_THREADS", threads)
        got[threads] = (graph.perm_apply(rp, ci, va, vomp), graph.order_rcm(rp, ci),
                        graph.renumber_rcm(rp, ci, va), graph.order_gorder(rp, ci, 3))
    for threads in ("5", "16"):
        for a, b in zip(got["1"], got[threads]):
            if isinstance(a, tuple):
                assert all(np.array_equal(x, y) for x, y in zip(a, b))
            else:
                assert np.array_equal(a, b)
    # the permuted matrix really is P·A·Pᵀ with sorted rows
    import scipy.sparse as sp
    prp, pci, pva = got["5"][0][:3]
    P = sp.csr_matrix((pva, pci, prp), shape=A.shape)
    want = A[vomp][:, vomp].tocsr()
    want.sort_indices()
    assert np.array_equal(P.indptr, want.indptr) and np.array_equal(P.indices, want.indices)
    assert np.array_equal(P.data, want.data)
    if R.have("ref_graph.so"):
        monkeypatch.setenv("GCNB_HOST_THREADS", "7")
        assert np.array_equal(graph.order_rcm(rp, ci), R.ref_order_rcm(rp, ci))

"""Parity of the CUDA SpMM path (through the C ABI) with the oracle and the reference's golden
answers.  Tolerance is BASELINE.json's north_star: 1e-5 relative for fp32, 1e-2 for bf16 features
(relative = max|x - ref| / max|ref|, ref accumulated in fp64)."""
import ctypes

import numpy as np
import pytest
import torch

from conftest import random_normalized_graph, rel_err
from gcn_b200 import _ffi, graph, ops, synth
from oracle import ref_bindings as R
from oracle import spmm_oracle as SO

pytestmark = pytest.mark.gpu
FP32_TOL = 1e-5
BF16_TOL = 1e-2
DEV = "cuda:0"


def plan_of(g, prefix="", **kw):
    rp, ci, va = g[prefix + "rowptr"], g[prefix + "col"], g[prefix + "val"]
    n = len(rp) - 1
    return ops.SpmmPlan(rp, ci, va, (n, n), device=DEV, **kw)


def test_golden_forward(golden_graph):
    g = golden_graph
    plan = plan_of(g)
    C = plan.spmm(torch.from_numpy(g["B"]).to(DEV)).cpu().numpy()
    assert rel_err(C, g["C_f64"]) < FP32_TOL
    assert rel_err(C, g["C_torch_spmm"]) < 2 * FP32_TOL
    assert plan.symmetric  # D^-1/2 (A+I) D^-1/2 of a symmetric pattern


def test_golden_transposed_matches_reference_autograd(golden_graph):
    g = golden_graph
    plan = plan_of(g, "U_")
    assert not plan.symmetric and plan.info()["has_transpose"] == 1
    Gt = torch.from_numpy(g["G"]).to(DEV)
    assert rel_err(plan.spmm_t(Gt).cpu().numpy(), g["UtG_autograd"]) < FP32_TOL
    # through autograd
    B = torch.from_numpy(g["B"]).to(DEV).requires_grad_(True)
    ops.spmm(plan, B).backward(Gt)
    assert rel_err(B.grad.cpu().numpy(), g["UtG_autograd"]) < FP32_TOL


@pytest.mark.parametrize("k", [1, 2, 3, 4, 7, 8, 16, 32, 47, 64, 100, 128, 200, 256, 512, 602])
def test_feature_width_sweep(k):
    A = random_normalized_graph(700, 6000, k, "skew")
    n = A.shape[0]
    rng = np.random.default_rng(k)
    B = rng.standard_normal((n, k)).astype(np.float32)
    plan = ops.SpmmPlan.from_scipy(A, device=DEV)
    C = plan.spmm(torch.from_numpy(B).to(DEV)).cpu().numpy()
    assert rel_err(C, SO.spmm_f64(A.indptr, A.indices, A.data, n, B)) < FP32_TOL


def test_long_rows_block_class_and_split_rows():
    # rows above kLongRow (512) go to the block class; rows above kChunk (2048) are split
    import scipy.sparse as sp
    n = 6000
    rng = np.random.default_rng(7)
    rows = np.concatenate((np.zeros(5000, dtype=np.int64), np.full(700, 17), rng.integers(0, n, 20000)))
    cols = np.concatenate((rng.choice(n, 5000, replace=False), rng.choice(n, 700, replace=False),
                           rng.integers(0, n, 20000)))
    a = sp.coo_matrix((np.ones(len(rows)), (rows, cols)), shape=(n, n)).tocsr()
    a = ((a + a.T) > 0).astype(np.float64)
    a.setdiag(0)
    a.eliminate_zeros()
    A = graph.normalize_adj(a)
    plan = ops.SpmmPlan.from_scipy(A, device=DEV)
    info = plan.info()
    assert info["rows_block_class"] >= 2 and info["rows_split"] >= 1 and info["max_row_nnz"] > 2048
    for k in (5, 32, 100, 256):
        B = rng.standard_normal((n, k)).astype(np.float32)
        ref = SO.spmm_f64(A.indptr, A.indices, A.data, n, B)
        Bd = torch.from_numpy(B).to(DEV)
        C1 = plan.spmm(Bd)
        assert rel_err(C1.cpu().numpy(), ref) < FP32_TOL
        assert torch.equal(C1, plan.spmm(Bd))  # no atomics: bit-reproducible run to run


def test_edge_cases():
    # empty matrix, empty rows, m % 8 != 0, k = 0, rectangular
    rp = np.zeros(6, dtype=np.int32)
    plan = ops.SpmmPlan(rp, np.zeros(0, np.int32), np.zeros(0, np.float32), (5, 5), device=DEV)
    C = plan.spmm(torch.ones(5, 3, device=DEV))
    assert torch.all(C == 0)
    assert plan.spmm(torch.ones(5, 0, device=DEV)).shape == (5, 0)
    import scipy.sparse as sp
    rng = np.random.default_rng(3)
    A = sp.random(13, 29, density=0.2, format="csr", dtype=np.float32, random_state=np.random.RandomState(3))
    A.sort_indices()
    B = rng.standard_normal((29, 6)).astype(np.float32)
    plan = ops.SpmmPlan.from_scipy(A, device=DEV)
    assert rel_err(plan.spmm(torch.from_numpy(B).to(DEV)).cpu().numpy(),
                   SO.spmm_f64(A.indptr, A.indices, A.data, 29, B)) < FP32_TOL
    G = rng.standard_normal((13, 6)).astype(np.float32)
    assert rel_err(plan.spmm_t(torch.from_numpy(G).to(DEV)).cpu().numpy(),
                   SO.spmm_t_f64(A.indptr, A.indices, A.data, 29, G)) < FP32_TOL


def test_argument_errors():
    A = random_normalized_graph(50, 200, 1)
    plan = ops.SpmmPlan.from_scipy(A, device=DEV)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        plan.spmm(torch.ones(50, 4))
    with pytest.raises(ValueError):
        plan.spmm(torch.ones(49, 4, device=DEV))
    with pytest.raises(TypeError):
        plan.spmm(torch.ones(50, 4, device=DEV, dtype=torch.float64))
    with pytest.raises(_ffi.GcnB200Error):  # unsorted columns
        ops.SpmmPlan(np.array([0, 2], np.int32), np.array([1, 0], np.int32), np.ones(2, np.float32),
                     (1, 2), device=DEV)
    fwd_only = ops.SpmmPlan.from_scipy(A, device=DEV, need_transpose=False)
    with pytest.raises(_ffi.GcnB200Error):
        fwd_only.spmm_t(torch.ones(50, 4, device=DEV))


def test_strided_operands_and_epilogues():
    A = random_normalized_graph(300, 2500, 5)
    n = A.shape[0]
    rng = np.random.default_rng(5)
    big = torch.from_numpy(rng.standard_normal((n, 80)).astype(np.float32)).to(DEV)
    plan = ops.SpmmPlan.from_scipy(A, device=DEV)
    for lo, hi in ((0, 32), (3, 40), (16, 80)):   # aligned, misaligned, wide views (ldb = 80)
        Bv = big[:, lo:hi]
        ref = SO.spmm_f64(A.indptr, A.indices, A.data, n, Bv.cpu().numpy())
        assert rel_err(plan.spmm(Bv).cpu().numpy(), ref) < FP32_TOL
        out = torch.zeros(n, 96, device=DEV)
        plan.spmm(Bv, out=out[:, 5:5 + hi - lo])
        assert rel_err(out[:, 5:5 + hi - lo].cpu().numpy(), ref) < FP32_TOL
        assert torch.all(out[:, :5] == 0) and torch.all(out[:, 5 + hi - lo:] == 0)
    B = big[:, :48].contiguous()
    ref = SO.spmm_f64(A.indptr, A.indices, A.data, n, B.cpu().numpy())
    bias = torch.from_numpy(rng.standard_normal(48).astype(np.float32)).to(DEV)
    got = plan.spmm(B, bias=bias, relu=True).cpu().numpy()
    assert rel_err(got, np.maximum(ref + bias.cpu().numpy().astype(np.float64), 0)) < FP32_TOL
    base = torch.from_numpy(rng.standard_normal((n, 48)).astype(np.float32)).to(DEV)
    acc = base.clone()
    plan.spmm(B, out=acc, accumulate=True)
    assert rel_err(acc.cpu().numpy(), ref + base.cpu().numpy()) < FP32_TOL


@pytest.mark.parametrize("k", [6, 16, 100, 256])
def test_bf16_features(k):
    A = random_normalized_graph(500, 5000, 11, "skew")
    n = A.shape[0]
    rng = np.random.default_rng(11)
    B = torch.from_numpy(rng.standard_normal((n, k)).astype(np.float32)).to(DEV)
    plan = ops.SpmmPlan.from_scipy(A, device=DEV)
    Bb = B.to(torch.bfloat16)
    C = plan.spmm(Bb).cpu().numpy()
    # exact in the bf16-rounded inputs (fp32 accumulate) ...
    assert rel_err(C, SO.spmm_f64(A.indptr, A.indices, A.data, n, Bb.float().cpu().numpy())) < FP32_TOL
    # ... and within the stated bf16 bound of the fp32-feature answer
    assert rel_err(C, SO.spmm_f64(A.indptr, A.indices, A.data, n, B.cpu().numpy())) < BF16_TOL


def test_host_buffer_entry_point(golden_graph):
    g = golden_graph
    plan = plan_of(g)
    C = plan.spmm_host(g["B"])
    assert rel_err(C, g["C_f64"]) < FP32_TOL
    up = plan_of(g, "U_")
    assert rel_err(up.spmm_host(g["G"], transpose=True), g["UtG_autograd"]) < FP32_TOL


def test_permute_rows_bit_exact(golden_graph):
    g = golden_graph
    X = torch.from_numpy(g["B"]).to(DEV)
    vomp = torch.from_numpy(g["gorder_vomp"]).to(DEV)
    got = ops.permute_rows(X, vomp).cpu().numpy()
    assert np.array_equal(got, SO.permute_rows(g["B"], g["gorder_vomp"]))


def test_reordered_pipeline_equivalence(golden_graph):
    """renumber -> permute X -> SpMM == SpMM then permute (gcn6.py:313-377 pipeline)."""
    g = golden_graph
    rp, ci, va, vomp = graph.renumber_gorder(g["rowptr"], g["col"], g["val"])
    n = len(rp) - 1
    plan = ops.SpmmPlan(rp, ci, va, (n, n), device=DEV)
    Xp = ops.permute_rows(torch.from_numpy(g["B"]).to(DEV), torch.from_numpy(vomp).to(DEV))
    C = plan.spmm(Xp).cpu().numpy()
    assert rel_err(C, g["C_f64"][vomp]) < FP32_TOL


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def test_compat_flexspmm_symbol_on_reference_tiles(golden_graph):
    """The reference's flexspmm() name and argument order (flexspmm.cu:499-502), fed the reference
    tile format, C pre-zeroed by the caller like pygcn/gcn6.py:37."""
    g = golden_graph
    n = len(g["rowptr"]) - 1
    t = graph.csr2tile(g["rowptr"], g["col"], g["val"], n_sm=148)
    k = g["B"].shape[1]
    d = {key: torch.from_numpy(np.ascontiguousarray(t[key])).to(DEV)
         for key in ("seg_rowPtr", "segNzCV", "segVoMap", "grouped_tailSeg", "next_seg")}
    B = torch.from_numpy(g["B"]).to(DEV)
    C = torch.zeros(n, k, device=DEV)
    torch.cuda.synchronize()
    _ffi.lib.flexspmm(_p(d["seg_rowPtr"]), _p(d["segNzCV"]), _p(d["segVoMap"]), _p(d["grouped_tailSeg"]),
                      _p(d["next_seg"]), n, n, k, t["n_segs"], _p(B), _p(C))
    torch.cuda.synchronize()
    assert rel_err(C.cpu().numpy(), g["C_f64"]) < FP32_TOL
    # cuspmm(): CSR SpMM with the reference's mistyped float* rowPtr (cuspmm.cu:23)
    rp = torch.from_numpy(g["rowptr"]).to(DEV)
    ci = torch.from_numpy(g["col"]).to(DEV)
    va = torch.from_numpy(g["val"]).to(DEV)
    C2 = torch.zeros(n, k, device=DEV)
    _ffi.lib.cuspmm(_p(rp), _p(ci), _p(va), _p(B), _p(C2), n, n, len(g["col"]), k)
    torch.cuda.synchronize()
    assert rel_err(C2.cpu().numpy(), g["C_f64"]) < FP32_TOL
    # permutate(): in place
    X = B.clone()
    vomp = torch.from_numpy(g["dfs_vomp"]).to(DEV)
    _ffi.lib.permutate(_p(X), _p(vomp), None, n, n, k)
    torch.cuda.synchronize()
    assert np.array_equal(X.cpu().numpy(), g["B"][g["dfs_vomp"]])


@pytest.mark.skipif(not R.have("flexspmm.so"), reason="compiled reference kernels not shipped")
def test_against_compiled_reference_kernel_on_gpu(golden_graph):
    """The reference's own flexspmm.so (unmodified flexspmm.cu built for sm_100a) on corrected
    reference tiles vs our engine, same inputs.  Also csr2tile itself (needs the device query)."""
    g = golden_graph
    n = len(g["rowptr"]) - 1
    k = g["B"].shape[1]
    t = R.ref_csr2tile_gpu(g["rowptr"], g["col"], g["val"], corrected=True)
    ours = graph.csr2tile(g["rowptr"], g["col"], g["val"],
                          n_sm=torch.cuda.get_device_properties(0).multi_processor_count)
    for key in ("n_segs", "segVoMap", "seg_rowPtr", "segNzCV"):
        assert np.array_equal(ours[key], t[key]), key
    n_sm = len(ours["next_seg"]) - 1
    assert np.array_equal(ours["next_seg"], t["next_seg"][: n_sm + 1])
    assert np.array_equal(ours["grouped_tailSeg"], t["grouped_tailSeg"][: n_sm + 1])
    lib = ctypes.CDLL(R.REF_DIR + "/flexspmm.so")
    d = {key: torch.from_numpy(np.ascontiguousarray(t[key])).to(DEV)
         for key in ("seg_rowPtr", "segNzCV", "segVoMap", "grouped_tailSeg", "next_seg")}
    B = torch.from_numpy(g["B"]).to(DEV)
    C = torch.zeros(n, k, device=DEV)
    torch.cuda.synchronize()
    lib.flexspmm(_p(d["seg_rowPtr"]), _p(d["segNzCV"]), _p(d["segVoMap"]), _p(d["grouped_tailSeg"]),
                 _p(d["next_seg"].clone()), n, n, k, t["n_segs"], _p(B), _p(C))
    torch.cuda.synchronize()
    ref = C.cpu().numpy()
    assert rel_err(ref, g["C_f64"]) < FP32_TOL          # the reference kernel itself is right on fixed tiles
    mine = plan_of(g).spmm(B).cpu().numpy()
    assert rel_err(mine, ref) < 2 * FP32_TOL


# ---- full-size properties (BASELINE shapes; the oracle would take too long) ----------------------
@pytest.mark.parametrize("shape,k", [("arxiv", 256), ("products", 100), ("reddit", 128), ("rmat", 64)])
def test_full_size_properties(shape, k):
    if shape == "rmat":     # R-MAT scale 22, edge factor 16: the sweep's shape at a size the test box generates fast
        rp, ci, va, n = synth.make_graph("rmat", device=DEV, scale=22, edge_factor=16)
    else:
        rp, ci, va, n = synth.make_graph(shape, device=DEV)
    if shape == "reddit":   # BASELINE config 3 names 114 M edges: the generator must realise them
        assert abs(len(ci) - 114_832_965) <= 1000
    plan = ops.SpmmPlan(rp, ci, va, (n, n), device=DEV)
    assert plan.symmetric
    if shape == "reddit":
        assert plan.info()["panel_items"] == 0    # the panel class is opt-in (measured: no gain)
    g = torch.Generator(device=DEV).manual_seed(1)
    X = torch.randn(n, k, device=DEV, generator=g)
    Y = torch.randn(n, k, device=DEV, generator=g)
    AX, AY = plan.spmm(X), plan.spmm(Y)
    # linearity: A(2X - 3Y) = 2AX - 3AY
    lin = plan.spmm(2 * X - 3 * Y)
    assert rel_err(lin.cpu().numpy(), (2 * AX - 3 * AY).cpu().numpy()) < 1e-5
    # checksum of checksums: 1ᵀ(AX) = (Aᵀ1)ᵀX, with Aᵀ1 from the host CSR in fp64
    colsum = np.bincount(ci, weights=va.astype(np.float64), minlength=n)
    lhs = AX.double().sum(0).cpu().numpy()
    rhs = colsum @ X.double().cpu().numpy()
    assert np.abs(lhs - rhs).max() / np.abs(rhs).max() < 1e-6
    # sampled rows against the oracle's arithmetic in fp64
    rows = np.random.default_rng(0).integers(0, n, 64)
    Xh = X.cpu().numpy().astype(np.float64)
    for r in rows:
        z = slice(rp[r], rp[r + 1])
        ref = va[z].astype(np.float64) @ Xh[ci[z]]
        assert np.abs(AX[r].cpu().numpy() - ref).max() <= 1e-5 * max(np.abs(ref).max(), 1e-3)
    # symmetric: <Y, AX> = <AY, X>
    a, b = (Y.double() * AX.double()).sum().item(), (AY.double() * X.double()).sum().item()
    assert abs(a - b) <= 1e-6 * max(abs(a), 1.0) + 1e-3


def _dense_skew_graph(n=1800, e=90000, seed=21):
    """Mean row length >= 24 (the packed stream kernel's dispatch condition, spmm.cu launch_vec) with
    a heavy head: some rows beyond kSegLong = 1024 non-zeros are cut into pieces, plus empty-ish tails."""
    A = random_normalized_graph(n, e, seed, "skew")
    assert A.nnz >= 24 * A.shape[0] and np.diff(A.indptr).max() > 1024
    return A


@pytest.mark.parametrize("k", [1, 2, 4, 8, 12, 16, 20, 32, 36, 47, 48, 64])
def test_packed_stream_kernel_narrow_widths(k):
    """Narrow feature rows on a graph dense enough for the packed stream kernel (S = 2/4/8 non-zeros
    per warp-wide gather): fp32 and bf16 operands, bias/ReLU epilogue, accumulate, bit-reproducible."""
    A = _dense_skew_graph()
    n = A.shape[0]
    rng = np.random.default_rng(100 + k)
    B = rng.standard_normal((n, k)).astype(np.float32)
    bias = rng.standard_normal(k).astype(np.float32)
    ref = SO.spmm_f64(A.indptr, A.indices, A.data, n, B)
    plan = ops.SpmmPlan.from_scipy(A, device=DEV)
    Bd = torch.from_numpy(B).to(DEV)
    C = plan.spmm(Bd)
    assert rel_err(C.cpu().numpy(), ref) < FP32_TOL
    assert torch.equal(C, plan.spmm(Bd))
    Ce = plan.spmm(Bd, bias=torch.from_numpy(bias).to(DEV), relu=True).cpu().numpy()
    assert rel_err(Ce, np.maximum(ref + bias, 0)) < FP32_TOL
    C0 = torch.from_numpy(rng.standard_normal((n, k)).astype(np.float32)).to(DEV)
    want = C0.cpu().numpy().astype(np.float64) + ref
    assert rel_err(plan.spmm(Bd, out=C0, accumulate=True).cpu().numpy(), want) < FP32_TOL
    Bb = Bd.to(torch.bfloat16)
    ref_b = SO.spmm_f64(A.indptr, A.indices, A.data, n, Bb.float().cpu().numpy())
    assert rel_err(plan.spmm(Bb).cpu().numpy(), ref_b) < FP32_TOL     # same bf16 inputs: fp32 accumulation
    assert rel_err(plan.spmm(Bb).cpu().numpy(), ref) < BF16_TOL
    # padded rows (row pitch > k) in and out
    Bp = torch.zeros(n, k + 4, device=DEV)[:, :k]
    Bp.copy_(Bd)
    Cp = torch.empty(n, k + 8, device=DEV)[:, :k]
    assert rel_err(plan.spmm(Bp, out=Cp).cpu().numpy(), ref) < FP32_TOL


@pytest.mark.parametrize("k", [4, 8, 12, 28, 32, 36, 60, 64, 68, 100, 104, 128, 132, 200, 256, 260, 512, 602])
@pytest.mark.parametrize("dense", [True, False])
def test_f32_rows_at_32_byte_pitch(k, dense):
    """fp32 feature rows whose pitch is a multiple of 8 floats (32-byte aligned rows): the shapes the 256-bit
    load kernels take (spmm.cu launch_vec8: packed S = 8/4/2 up to k = 128 on dense-row graphs, stream NV = 1/2
    beyond; k % 8 == 4 leaves the last lane 4 real and 4 padding columns).  Padding columns of C must stay
    untouched, epilogues and accumulate as everywhere else, bit-reproducible."""
    A = _dense_skew_graph() if dense else random_normalized_graph(900, 9000, k, "skew")
    n = A.shape[0]
    rng = np.random.default_rng(500 + k)
    B = rng.standard_normal((n, k)).astype(np.float32)
    bias = rng.standard_normal(k).astype(np.float32)
    ref = SO.spmm_f64(A.indptr, A.indices, A.data, n, B)
    plan = ops.SpmmPlan.from_scipy(A, device=DEV)
    pitch = (k + 7) // 8 * 8
    for extra in (0, 8):
        Bfull = torch.full((n, pitch + extra), 1e30, device=DEV)      # poison in the padding: must never reach C
        Bp = Bfull[:, :k]
        Bp.copy_(torch.from_numpy(B))
        Cfull = torch.full((n, pitch + extra), -7.0, device=DEV)
        Cp = Cfull[:, :k]
        out = plan.spmm(Bp, out=Cp)
        assert rel_err(out.cpu().numpy(), ref) < FP32_TOL
        assert bool((Cfull[:, k:] == -7.0).all()), "padding columns of C were written"
        again = torch.full((n, pitch + extra), -7.0, device=DEV)[:, :k]
        assert torch.equal(plan.spmm(Bp, out=again), out)
        Ce = torch.empty(n, pitch + extra, device=DEV)[:, :k]
        plan.spmm(Bp, out=Ce, bias=torch.from_numpy(bias).to(DEV), relu=True)
        assert rel_err(Ce.cpu().numpy(), np.maximum(ref + bias, 0)) < FP32_TOL
        C0 = torch.zeros(n, pitch + extra, device=DEV)[:, :k]
        C0.copy_(torch.from_numpy(B))
        want = B.astype(np.float64) + ref
        assert rel_err(plan.spmm(Bp, out=C0, accumulate=True).cpu().numpy(), want) < FP32_TOL


# ---- panel class (32-row panels walking their columns in step; spmm_panel_kernel) -----------------
def _hub_graph(n, e, seed):
    """Skewed rows AND columns: the first panels are nearly dense (they get cut into several items whose
    partial rows meet in the fixed-order reduce), the tail rows hold a few non-zeros."""
    import scipy.sparse as sp
    rng = np.random.default_rng(seed)
    r = (n * rng.random(e) ** 3).astype(np.int64)
    c = (n * rng.random(e) ** 2).astype(np.int64)
    a = sp.coo_matrix((np.ones(e), (r, c)), shape=(n, n)).tocsr()
    a = ((a + a.T) > 0).astype(np.float64)
    a.setdiag(0)
    a.eliminate_zeros()
    return graph.normalize_adj(a)


@pytest.mark.parametrize("n,e", [(4001, 1_500_000), (700, 6000), (2500, 300_000)])
@pytest.mark.parametrize("k", [68, 100, 128, 200, 602])
def test_panel_class_matches_oracle(n, e, k):
    A = _hub_graph(n, e, k + n)
    rng = np.random.default_rng(k)
    B = rng.standard_normal((n, k)).astype(np.float32)
    plan = ops.SpmmPlan.from_scipy(A, device=DEV, panel=True)
    info = plan.info()
    assert info["panel_items"] >= (n + 31) // 32 and info["panel_distinct_columns"] > 0
    Bd = torch.from_numpy(B).to(DEV)
    C = plan.spmm(Bd)
    ref = SO.spmm_f64(A.indptr, A.indices, A.data, n, B)
    assert rel_err(C.cpu().numpy(), ref) < FP32_TOL
    # no atomics anywhere: bit-reproducible, also against a second plan
    assert torch.equal(C, plan.spmm(Bd))
    assert torch.equal(C, ops.SpmmPlan.from_scipy(A, device=DEV, panel=True).spmm(Bd))
    # the stream class on the same matrix agrees to rounding
    C2 = ops.SpmmPlan.from_scipy(A, device=DEV, panel=False).spmm(Bd)
    assert rel_err(C2.cpu().numpy(), ref) < FP32_TOL
    # epilogue flags and transposed product (symmetric: same plan)
    bias = torch.from_numpy(rng.standard_normal(k).astype(np.float32)).to(DEV)
    Cb = plan.spmm(Bd, bias=bias, relu=True).cpu().numpy()
    assert rel_err(Cb, np.maximum(ref + bias.cpu().numpy(), 0)) < FP32_TOL
    assert rel_err(plan.spmm_t(Bd).cpu().numpy(), SO.spmm_t_f64(A.indptr, A.indices, A.data, n, B)) < FP32_TOL


def test_panel_class_items_are_cut_and_reduced():
    """The dense head of _hub_graph(4001, 1.5 M) exceeds 32 k non-zeros per panel: several items per panel."""
    A = _hub_graph(4001, 1_500_000, 5)
    assert int(A.indptr[32]) > 2 * 32768
    plan = ops.SpmmPlan.from_scipy(A, device=DEV, panel=True)
    assert plan.info()["panel_items"] > (4001 + 31) // 32


@pytest.mark.parametrize("k", [100, 256])
def test_panel_class_bf16_and_strides(k):
    A = _hub_graph(2500, 300_000, 9)
    n = A.shape[0]
    rng = np.random.default_rng(3)
    B = rng.standard_normal((n, k)).astype(np.float32)
    plan = ops.SpmmPlan.from_scipy(A, device=DEV, panel=True)
    Bb = torch.from_numpy(B).to(DEV).to(torch.bfloat16)
    ref = SO.spmm_f64(A.indptr, A.indices, A.data, n, Bb.float().cpu().numpy())
    assert rel_err(plan.spmm(Bb).cpu().numpy(), ref) < 1e-5      # exact bf16 inputs, fp32 accumulation
    assert rel_err(plan.spmm(Bb).cpu().numpy(), SO.spmm_f64(A.indptr, A.indices, A.data, n, B)) < BF16_TOL
    # padded pitch of B and of C
    Bp = torch.zeros(n, k + 12, device=DEV)
    Bp[:, :k] = torch.from_numpy(B).to(DEV)
    out = torch.full((n, k + 8), 7.0, device=DEV)
    plan.spmm(Bp[:, :k], out=out[:, :k])
    assert rel_err(out[:, :k].cpu().numpy(), SO.spmm_f64(A.indptr, A.indices, A.data, n, B)) < FP32_TOL
    assert torch.all(out[:, k:] == 7.0)

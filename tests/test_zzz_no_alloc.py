"""gcnb_colsum_noalloc: the column sums of G through a static device scratch array, no allocation at all.

The entry point has no counterpart in the reference (it was added for callers
that cannot call cudaMalloc); its check is the existing entry point gcnb_xw_tn_colsum (M = 0: the two-pass column sums
of the same kernels with a pool allocation) plus an fp64 sum.  Collected last (file name) so that a failure here
hides nothing under ``pytest -x``.
"""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None else 0)


def _noalloc(ffi, G, out, K=None, N=None):
    K = G.shape[0] if K is None else K
    N = G.shape[1] if N is None else N
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    return ffi.lib.gcnb_colsum_noalloc(_ptr(G), G.stride(0), K, N, _ptr(out), stream)


def _existing(ffi, G, out):
    K, N = G.shape
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ffi.call("gcnb_xw_tn_colsum", None, 0, _ptr(G), G.stride(0), None, N, _ptr(out), 0, N, K, 0, stream)


@pytest.mark.gpu
def test_colsum_noalloc_matches_existing_entry_point_twice():
    from gcn_b200 import _ffi
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(1234)
    # two calls in a row with different shapes: the second finds the first call's partial sums in the scratch
    shapes = [(20_000, 47, 64), (3_001, 100, 100)]
    outs, refs, mats = [], [], []
    for K, N, ld in shapes:
        G = torch.randn(K, ld, generator=g, device=dev, dtype=torch.float32)[:, :N]
        out = torch.full((N,), float("nan"), device=dev)
        rc = _noalloc(_ffi, G, out)
        assert rc == 0, _ffi.last_error()
        outs.append(out)
        mats.append(G)
    for G in mats:
        ref = torch.empty(G.shape[1], device=dev)
        _existing(_ffi, G, ref)
        refs.append(ref)
    torch.cuda.synchronize()
    for G, out, ref in zip(mats, outs, refs):
        assert torch.equal(out, ref), "same kernels, same order: the two entry points agree bit for bit"
        exact = G.double().sum(0)
        scale = G.double().abs().sum(0)
        assert float(((out.double() - exact).abs() / scale).max()) <= 1e-5
    # same input again: the result does not depend on what the scratch held
    again = torch.empty_like(outs[0])
    assert _noalloc(_ffi, mats[0], again) == 0, _ffi.last_error()
    torch.cuda.synchronize()
    assert torch.equal(again, outs[0])


@pytest.mark.gpu
def test_colsum_noalloc_allocates_nothing_and_rejects_oversize():
    from gcn_b200 import _ffi
    dev = torch.device("cuda:0")
    G = torch.randn(4096, 32, device=dev)
    out = torch.empty(32, device=dev)
    assert _noalloc(_ffi, G, out) == 0, _ffi.last_error()     # first call looks the symbol address up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert _noalloc(_ffi, G, out) == 0, _ffi.last_error()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, "the call changed the device's free memory"
    np.testing.assert_allclose(out.cpu().numpy(), G.double().sum(0).cpu().numpy(), rtol=0, atol=1e-5 * 4096)
    # empty and degenerate sizes
    z = torch.full((32,), 7.0, device=dev)
    assert _noalloc(_ffi, G, z, K=0) == 0
    torch.cuda.synchronize()
    assert float(z.abs().max()) == 0.0
    assert _noalloc(_ffi, G, z, N=0) == 0
    # beyond the static scratch: a clear error, nothing launched
    cap = int(_ffi.lib.gcnb_colsum_noalloc_max_floats())
    assert cap == 2 << 20
    K_big = (cap // 32 + 1) * 512
    rc = _noalloc(_ffi, G, out, K=K_big, N=32)
    assert rc != 0
    assert "static scratch" in _ffi.last_error()
    # still usable afterwards
    assert _noalloc(_ffi, G, out) == 0, _ffi.last_error()
    torch.cuda.synchronize()

"""gcnb_xw (tcgen05 tensor-core X·W) against an fp64 torch reference of the same op."""
import numpy as np
import pytest
import torch

from gcn_b200 import gemm

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = {0: 1e-5, 1: 3e-3, 2: 2e-2}   # 3xTF32 (fp32 parity, north_star 1e-5) / TF32 / bf16 operands


def rel(x, ref):
    return float((x.double() - ref).abs().max() / ref.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("M,N,K", [(128, 16, 16), (300, 7, 33), (2708, 16, 1433), (1000, 256, 100),
                                   (517, 47, 256), (4096, 128, 128), (129, 130, 17), (5, 3, 2)])
@pytest.mark.parametrize("precision", [0, 1, 2])
def test_xw_forward(M, N, K, precision):
    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    X = torch.randn(M, K, device=DEV, generator=g)
    W = torch.randn(K, N, device=DEV, generator=g) / K ** 0.5
    out = gemm.gemm_raw(X, W, precision=precision)
    ref = X.double() @ W.double()
    assert rel(out, ref) < TOL[precision]


def test_xw_transposes_bias_relu_and_views():
    g = torch.Generator(device=DEV).manual_seed(3)
    M, N, K = 777, 40, 100
    X = torch.randn(M, K, device=DEV, generator=g)
    W = torch.randn(K, N, device=DEV, generator=g)
    b = torch.randn(N, device=DEV, generator=g)
    ref = X.double() @ W.double()
    assert rel(gemm.gemm_raw(X.t().contiguous(), W, trans_a=True), ref) < 1e-5
    assert rel(gemm.gemm_raw(X, W.t().contiguous(), trans_b=True), ref) < 1e-5
    assert rel(gemm.gemm_raw(X.t().contiguous(), W.t().contiguous(), trans_a=True, trans_b=True), ref) < 1e-5
    got = gemm.gemm_raw(X, W, bias=b, relu=True)
    assert rel(got, torch.relu(ref + b.double())) < 1e-5
    # strided views: leading dimensions larger than the row, misaligned bases (scalar copy path)
    big = torch.randn(M, K + 9, device=DEV, generator=g)
    Xv = big[:, 3:3 + K]
    assert rel(gemm.gemm_raw(Xv, W), Xv.double() @ W.double()) < 1e-5
    out = torch.zeros(M, N + 5, device=DEV)
    gemm.gemm_raw(X, W, out=out[:, 2:2 + N])
    assert rel(out[:, 2:2 + N], ref) < 1e-5 and torch.all(out[:, :2] == 0) and torch.all(out[:, 2 + N:] == 0)


def test_xw_tall_k_split_and_autograd():
    """dW = Xᵀ·dS has a huge inner dimension: exercises split-K; and the autograd wiring."""
    g = torch.Generator(device=DEV).manual_seed(5)
    n, fin, fout = 50_000, 100, 64
    X = torch.randn(n, fin, device=DEV, generator=g)
    dS = torch.randn(n, fout, device=DEV, generator=g)
    ref = X.double().t() @ dS.double()
    assert rel(gemm.gemm_raw(X, dS, trans_a=True), ref) < 1e-5
    Xs = X[:3000].clone().requires_grad_(True)
    W = (torch.randn(fin, fout, device=DEV, generator=g) / 10).requires_grad_(True)
    b = torch.randn(fout, device=DEV, generator=g).requires_grad_(True)
    out = gemm.xw(Xs, W, bias=b, relu=True)
    G = torch.randn_like(out)
    out.backward(G)
    Xr, Wr, br = (t.detach().double().requires_grad_(True) for t in (Xs, W, b))
    torch.relu(Xr @ Wr + br).backward(G.double())
    assert rel(Xs.grad, Xr.grad) < 1e-5 and rel(W.grad, Wr.grad) < 1e-5 and rel(b.grad, br.grad) < 1e-5


@pytest.mark.parametrize("K,M,N", [(20_000, 256, 100), (20_000, 100, 256), (9_000, 256, 47), (5_000, 47, 256),
                                   (4_100, 16, 7), (33_333, 300, 130), (4_096, 128, 128)])
@pytest.mark.parametrize("precision", [0, 1, 2])
def test_xw_tall_k_raw_second_operand(K, M, N, precision, monkeypatch):
    """Tall-K variant (GCNB_XW_TALLK=1): the second operand of dW = Xᵀ·dS is converted inside the kernel
    instead of through the prepared image.  Same numbers as the image path, both against fp64.
    The TMEM accumulator does not round to nearest, so the error grows with the MMAs chained into one
    split (xw_tcgen05.cu bounds the chain); the longest K gets twice the fp32 tolerance."""
    tol = TOL[precision] * (2 if K > 30_000 else 1)
    g = torch.Generator(device=DEV).manual_seed(K + M + N)
    A = torch.randn(K, M, device=DEV, generator=g)
    wide = torch.randn(K, N + 1, device=DEV, generator=g)
    ref_of = lambda B: A.double().t() @ B.double()
    for B in (wide[:, :N].contiguous(),      # leading dimension N (scalar copies when N % 4 != 0)
              wide[:, :N],                   # a view: leading dimension N + 1
              wide[:, 1:]):                  # a view whose base is 4-byte aligned only
        ref = ref_of(B)
        monkeypatch.setenv("GCNB_XW_TALLK", "1")
        got = gemm.gemm_raw(A, B, trans_a=True, precision=precision)
        monkeypatch.setenv("GCNB_XW_TALLK", "0")
        img = gemm.gemm_raw(A, B, trans_a=True, precision=precision)
        assert rel(got, ref) < tol, (K, M, N, precision, tuple(B.stride()))
        assert rel(img, ref) < tol
        assert rel(got, img.double()) < 2 * tol
    # K-contiguous first operand (no transpose), tall K all the same
    At = A.t().contiguous()
    monkeypatch.setenv("GCNB_XW_TALLK", "1")
    assert rel(gemm.gemm_raw(At, wide[:, :N].contiguous(), precision=precision), ref_of(wide[:, :N])) < tol


def test_xw_degenerate():
    X = torch.randn(10, 0, device=DEV)
    W = torch.randn(0, 4, device=DEV)
    b = torch.arange(4, device=DEV, dtype=torch.float32) - 1
    assert torch.equal(gemm.gemm_raw(X, W, bias=b, relu=True), torch.relu(b).expand(10, 4))
    assert gemm.gemm_raw(torch.randn(0, 3, device=DEV), torch.randn(3, 4, device=DEV)).shape == (0, 4)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        gemm.gemm_raw(torch.randn(2, 2), torch.randn(2, 2))

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


# ---- streaming kernel (xw_stream.cu): persistent warp-specialised tcgen05 kernel for M >> K, N --------------
@pytest.mark.parametrize("M,N,K", [(2500, 256, 100), (100_000, 48, 256), (5000, 256, 48), (1024, 64, 32),
                                   (3001, 47, 100), (1500, 130, 260), (20_000, 300, 36), (1100, 8, 4)])
@pytest.mark.parametrize("precision", [0, 1])
def test_xw_stream_shapes(M, N, K, precision, monkeypatch):
    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    X = torch.randn(M, K, device=DEV, generator=g)
    W = torch.randn(K, N, device=DEV, generator=g) / K ** 0.5
    b = torch.randn(N, device=DEV, generator=g)
    ref = X.double() @ W.double()
    out = gemm.gemm_raw(X, W, precision=precision)
    assert rel(out, ref) < TOL[precision]
    assert rel(gemm.gemm_raw(X, W.t().contiguous(), trans_b=True, precision=precision), ref) < TOL[precision]
    assert rel(gemm.gemm_raw(X, W, bias=b, relu=True, precision=precision), torch.relu(ref + b.double())) < TOL[precision]
    # the first kernel on the same operands (GCNB_XW_STREAM is read once per process: compare through the
    # transposed-X form, which always takes the first kernel)
    old = gemm.gemm_raw(X.t().contiguous(), W, trans_a=True, precision=precision)
    assert rel(old, ref) < TOL[precision]
    # padded leading dimensions of X and D
    big = torch.randn(M, K + 12, device=DEV, generator=g)
    Xv = big[:, 4:4 + K]
    dst = torch.full((M, N + 8), 3.0, device=DEV)
    gemm.gemm_raw(Xv, W, out=dst[:, 4:4 + N] if N % 4 == 0 else dst[:, 1:1 + N], precision=precision)
    got = dst[:, 4:4 + N] if N % 4 == 0 else dst[:, 1:1 + N]
    assert rel(got, Xv.double() @ W.double()) < TOL[precision]
    assert torch.all(dst[:, 0] == 3.0) and torch.all(dst[:, -1] == 3.0)


def test_xw_fused_dropout_and_mask():
    g = torch.Generator(device=DEV).manual_seed(11)
    M, N, K = 1500, 24, 20
    X = torch.randn(M, K, device=DEV, generator=g)
    W = torch.randn(K, N, device=DEV, generator=g)
    b = torch.randn(N, device=DEV, generator=g)
    p, seed = 0.375, 12345     # multiples of 1/256 are exact (the kernel quantises p to 1/256)
    keep = gemm.dropout_keep_mask(M, N, p, seed).to(DEV)
    assert 0.58 < keep.float().mean().item() < 0.67
    ref = torch.relu(X.double() @ W.double() + b.double()) * keep / (1 - p)
    got = gemm.gemm_raw(X, W, bias=b, relu=True, dropout=(p, seed))
    assert rel(got, ref) < 1e-5
    assert not torch.equal(got, gemm.gemm_raw(X, W, bias=b, relu=True, dropout=(p, seed + 1)))   # another step, another mask
    # the tail kernel of the first GEMM kernel draws the same mask (transposed-X form of the same product)
    got_t = gemm.gemm_raw(X.t().contiguous(), W, trans_a=True, bias=b, relu=True, dropout=(p, seed))
    assert rel(got_t, ref) < 1e-5
    # the bit mask [out > 0] written next to the product, on both kernels
    bits = torch.empty(M, (N + 31) // 32, dtype=torch.int32, device=DEV)
    got2 = gemm.gemm_raw(X, W, bias=b, relu=True, dropout=(p, seed), mask_out=bits)
    assert torch.equal(got2, got) and torch.equal(gemm.unpack_mask_bits(bits, N), got > 0)
    bits_t = torch.empty_like(bits)
    gemm.gemm_raw(X.t().contiguous(), W, trans_a=True, bias=b, relu=True, dropout=(p, seed), mask_out=bits_t)
    assert torch.equal(gemm.unpack_mask_bits(bits_t, N), got_t > 0)
    # mask epilogue: (G·W2) ⊙ bits / (1-p)  (G: M x K2, product M x N)
    K2 = 40
    G = torch.randn(M, K2, device=DEV, generator=g)
    W2 = torch.randn(K2, N, device=DEV, generator=g)
    refm = (G.double() @ W2.double()) * (got > 0) / (1 - p)
    assert rel(gemm.gemm_raw(G, W2, mask=(bits, p)), refm) < 1e-5
    assert rel(gemm.gemm_raw(G.t().contiguous(), W2, trans_a=True, mask=(bits, p)), refm) < 1e-5


@pytest.mark.parametrize("p", [0.0, 0.5])
def test_mlp2_block_matches_unfused_autograd(p):
    """gemm.mlp2 (fused dense block of the 2-layer GCN) against torch autograd in fp64 with the SAME dropout mask."""
    g = torch.Generator(device=DEV).manual_seed(17)
    M, K, H, N = 4000, 36, 64, 12
    X = torch.randn(M, K, device=DEV, generator=g)
    W1 = (torch.randn(K, H, device=DEV, generator=g) / 6).requires_grad_(True)
    b1 = torch.randn(H, device=DEV, generator=g).requires_grad_(True)
    W2 = (torch.randn(H, N, device=DEV, generator=g) / 8).requires_grad_(True)
    seed = 99
    s2 = gemm.mlp2(X, W1, b1, W2, p=p, seed=seed, training=True)
    G = torch.randn(M, N, device=DEV, generator=g)
    s2.backward(G)
    keep = (gemm.dropout_keep_mask(M, H, p, seed).to(DEV) if p > 0 else torch.ones(M, H, dtype=torch.bool, device=DEV))
    W1r, b1r, W2r = (t.detach().double().requires_grad_(True) for t in (W1, b1, W2))
    hr = torch.relu(X.double() @ W1r + b1r) * keep / (1 - p)
    (hr @ W2r).backward(G.double())
    assert rel(s2.detach(), (hr @ W2r).detach()) < 1e-5
    assert rel(W1.grad, W1r.grad) < 2e-5 and rel(b1.grad, b1r.grad) < 2e-5 and rel(W2.grad, W2r.grad) < 2e-5
    # eval mode: no dropout
    s_eval = gemm.mlp2(X, W1, b1, W2, p=p, seed=seed, training=False)
    assert rel(s_eval.detach(), (torch.relu(X.double() @ W1r + b1r) @ W2r).detach()) < 1e-5


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two visible GPUs")
def test_xw_on_second_device_after_first():
    """The > 48 KB shared-memory opt-in of the tensor-core kernels is a per-device function attribute:
    every kernel family on cuda:0 first, then on cuda:1, in one process."""
    for dev in ("cuda:0", "cuda:1"):
        g = torch.Generator(device=dev).manual_seed(4)
        for (M, N, K, ta) in [(300, 40, 100, False), (2048, 96, 64, False), (5000, 200, 36, False),
                              (64, 48, 6000, True), (100, 256, 8192, True)]:
            X = torch.randn(K, M, device=dev, generator=g) if ta else torch.randn(M, K, device=dev, generator=g)
            W = torch.randn(K, N, device=dev, generator=g) / K ** 0.5
            ref = (X.double().t() if ta else X.double()) @ W.double()
            assert rel(gemm.gemm_raw(X, W, trans_a=ta), ref) < 1e-5, (dev, M, N, K, ta)


@pytest.mark.parametrize("K,M,N", [(400_000, 100, 256), (300_000, 256, 48), (70_001, 128, 128), (5_000, 4, 8),
                                   (123_457, 48, 200)])
@pytest.mark.parametrize("precision", [0, 1])
def test_xw_tall_k_streaming_kernel(K, M, N, precision):
    """dW = Xᵀ·G on the MN-major streaming kernel (xw_tallk_stream.cu): several accumulation chains per CTA
    (K / 148 / 16 > 64 stages), the transposed route for M > 128, ragged K, padded leading dimensions; against
    fp64 and against the first kernel's tall-K variant."""
    g = torch.Generator(device=DEV).manual_seed(K % 1000 + M + N)
    X = torch.randn(K, M, device=DEV, generator=g)
    G = torch.randn(K, N, device=DEV, generator=g)
    ref = X.double().t() @ G.double()
    tol = TOL[precision] * (2 if K > 100_000 else 1)
    assert rel(gemm.gemm_raw(X, G, trans_a=True, precision=precision), ref) < tol
    # padded leading dimensions (views), bias + ReLU in the reduction
    Xp = torch.randn(K, M + 8, device=DEV, generator=g)
    Gp = torch.randn(K, N + 4, device=DEV, generator=g)
    b = torch.randn(N, device=DEV, generator=g)
    got = gemm.gemm_raw(Xp[:, 4:4 + M], Gp[:, :N], trans_a=True, bias=b, relu=True, precision=precision)
    refp = torch.relu(Xp[:, 4:4 + M].double().t() @ Gp[:, :N].double() + b.double())
    assert rel(got, refp) < tol


@pytest.mark.parametrize("K,M,N", [(400_000, 100, 256), (70_001, 48, 64), (50_000, 128, 64), (300_000, 256, 48),
                                   (3_000, 100, 256), (40_000, 30, 50), (9_000, 16, 7)])
def test_dw_colsum_one_pass(K, M, N):
    """gcnb_xw_tn_colsum: dW = Xᵀ·G and db = 1ᵀ·G from one pass (ones column in the 128-row tile of the tall-K
    streaming kernel when M < 128; two fixed-order passes over G otherwise — M = 128 / 256, small K, widths that
    are not multiples of 4).  Both against fp64; repeated calls are bit-identical."""
    g = torch.Generator(device=DEV).manual_seed(K % 997 + M + N)
    X = torch.randn(K, M, device=DEV, generator=g)
    G = torch.randn(K, N, device=DEV, generator=g) + 0.25      # a non-zero mean: the sums do not cancel
    dW, db = gemm.dw_colsum(X, G)
    tol = TOL[0] * (2 if K > 100_000 else 1)
    assert rel(dW, X.double().t() @ G.double()) < tol
    assert rel(db, G.double().sum(0)) < 1e-5     # north_star's fp32 tolerance; torch's own fp32 sum: ~1e-6
    dW2, db2 = gemm.dw_colsum(X, G)
    assert torch.equal(dW, dW2) and torch.equal(db, db2)
    # padded leading dimensions
    Xp = torch.randn(K, M + 8, device=DEV, generator=g)
    Gp = torch.randn(K, N + 4, device=DEV, generator=g)
    dW, db = gemm.dw_colsum(Xp[:, 4:4 + M], Gp[:, :N])
    assert rel(dW, Xp[:, 4:4 + M].double().t() @ Gp[:, :N].double()) < tol
    assert rel(db, Gp[:, :N].double().sum(0)) < 1e-5


def test_xw_autograd_bias_gradient_one_pass():
    """gemm.xw with a bias: weight and bias gradients come out of gcnb_xw_tn_colsum and match torch autograd."""
    g = torch.Generator(device=DEV).manual_seed(5)
    X = torch.randn(60_000, 100, device=DEV, generator=g)
    W = torch.randn(100, 64, device=DEV, generator=g, requires_grad=True)
    b = torch.randn(64, device=DEV, generator=g, requires_grad=True)
    go = torch.randn(60_000, 64, device=DEV, generator=g)
    gemm.xw(X, W, b).backward(go)
    gW, gb = W.grad.clone(), b.grad.clone()
    assert rel(gW, X.double().t() @ go.double()) < 2 * TOL[0]
    assert rel(gb, go.double().sum(0)) < 1e-5

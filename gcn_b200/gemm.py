"""Dense feature transform ``X·W (+bias)(ReLU)`` of the GCN layer — replaces ``torch.mm(input,
self.weight)`` (pygcn/gcn5.py:88, :136; gcn6.py:131, :191).

Forward and both backward products run on the tcgen05 kernel of libgcn_b200.so (``gcnb_xw``,
gcn_b200/csrc/xw_tcgen05.cu): fp32 operands are split on the fly into TF32 hi/lo parts in shared
memory and accumulated with three tensor-core passes into a TMEM accumulator, which keeps the
result within fp32 rounding of ``torch.mm`` (precision 0); precision 1 is single-pass TF32,
precision 2 bf16 operands.
"""
from __future__ import annotations

import ctypes
import os

import torch

from . import _ffi

PRECISION_3XTF32, PRECISION_TF32, PRECISION_BF16 = 0, 1, 2
_SWAP_DW = os.environ.get("GCNB_XW_SWAP_DW", "1") != "0"


def _row_major(t: torch.Tensor) -> torch.Tensor:
    if t.stride(1) != 1 or t.stride(0) < t.shape[1]:
        t = t.contiguous()
    return t


def gemm_raw(A: torch.Tensor, B: torch.Tensor, trans_a=False, trans_b=False, bias=None, relu=False,
             precision=PRECISION_3XTF32, out=None, dropout=None, mask=None, mask_out=None) -> torch.Tensor:
    """D = op(A)·op(B) (+bias)(ReLU)(dropout)(mask) on tensor cores, no autograd.  A, B float32 CUDA
    matrices (row-major, any leading dimension); op = transpose when the flag is set.
    dropout = (p, seed): after bias/ReLU, D = keep ? D / (1-p) : 0, keep bits from a counter hash, p quantised
    to 1/256 (gcnb_xw_ex, GCNB_EPI_DROPOUT).  mask_out: int32 tensor [M, ceil(N/32)] receiving the bits
    [D > 0].  mask = (bits, p): D = bit ? D / (1-p) : 0 (GCNB_EPI_MASK) — the backward of ReLU followed by
    dropout, with the bits its forward product wrote."""
    for t, nm in ((A, "A"), (B, "B")):
        if not t.is_cuda:
            raise RuntimeError(f"gcn_b200: GEMM operand {nm} is on '{t.device}'; CUDA only, no CPU fallback")
        if t.dtype != torch.float32 or t.dim() != 2:
            raise TypeError(f"gcn_b200: GEMM operand {nm} must be a 2-D float32 tensor")
    A = _row_major(A)
    B = _row_major(B)
    M, K = (A.shape[1], A.shape[0]) if trans_a else (A.shape[0], A.shape[1])
    K2, N = (B.shape[1], B.shape[0]) if trans_b else (B.shape[0], B.shape[1])
    if K != K2:
        raise ValueError(f"inner dimensions differ: {K} vs {K2}")
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32, device=A.device)
    epi = 0
    bias_ptr = None
    if bias is not None:
        bias = bias.contiguous()
        if bias.numel() != N or bias.dtype != torch.float32:
            raise ValueError("bias must be a float32 vector of length N")
        bias_ptr = ctypes.c_void_p(bias.data_ptr())
        epi |= _ffi.EPI_BIAS
    if relu:
        epi |= _ffi.EPI_RELU
    words = (N + 31) // 32

    def _bits(t, what):
        if t.dtype != torch.int32 or tuple(t.shape) != (M, words) or not t.is_contiguous() or t.device != A.device:
            raise ValueError(f"{what} must be a contiguous int32 tensor of shape (M, ceil(N/32)) on the operands' device")
        return ctypes.c_void_p(t.data_ptr())

    min_ptr = mout_ptr = None
    p_drop, seed = 0.0, 0
    if dropout is not None:
        p_drop, seed = float(dropout[0]), int(dropout[1]) & 0xFFFFFFFF
        if p_drop > 0.0:
            epi |= _ffi.EPI_DROPOUT
    if mask is not None:
        bits, p_mask = mask
        if dropout is not None and p_drop > 0.0 and float(p_mask) != p_drop:
            raise ValueError("dropout and mask share one scale 1/(1-p)")
        min_ptr, p_drop = _bits(bits, "mask bits"), float(p_mask)
        epi |= _ffi.EPI_MASK
    if mask_out is not None:
        mout_ptr = _bits(mask_out, "mask_out")
    stream = ctypes.c_void_p(torch.cuda.current_stream(A.device).cuda_stream)
    with torch.cuda.device(A.device):
        _ffi.call("gcnb_xw_ex", ctypes.c_void_p(A.data_ptr()), A.stride(0), int(trans_a),
                  ctypes.c_void_p(B.data_ptr()), B.stride(0), int(trans_b),
                  ctypes.c_void_p(out.data_ptr()), out.stride(0), M, N, K, bias_ptr, epi,
                  min_ptr, mout_ptr, words, p_drop, seed, int(precision), stream)
    return out


def unpack_mask_bits(bits: torch.Tensor, N: int) -> torch.Tensor:
    """[M, ceil(N/32)] int32 words -> [M, N] bool (bit c % 32 of word c // 32)."""
    M = bits.shape[0]
    sh = torch.arange(32, device=bits.device, dtype=torch.int64)
    b = ((bits.to(torch.int64).unsqueeze(-1) >> sh) & 1).bool().reshape(M, -1)
    return b[:, :N]


def dropout_keep_mask(M: int, N: int, p: float, seed: int) -> torch.Tensor:
    """The keep mask GCNB_EPI_DROPOUT draws for an M x N output (host restatement through
    gcnb_dropout_keeps, element by element — for tests on small shapes)."""
    out = torch.empty((M, N), dtype=torch.bool)
    for r in range(M):
        for c in range(N):
            out[r, c] = bool(_ffi.lib.gcnb_dropout_keeps(r, c, N, p, seed & 0xFFFFFFFF))
    return out


class _MLP2(torch.autograd.Function):
    """s2 = dropout(relu(X·W1 + b1)) · W2 — the dense part of the 2-layer GCN between the two aggregations
    when both layers run in the min-width association (gcn_b200.model), as ONE autograd node so that the
    elementwise neighbours of the products ride in their epilogues:
      forward   h  = xw_ex(X, W1, b1, ReLU | DROPOUT)          (replaces mm, +bias, relu, F.dropout: gcn6.py:131,142,245-246)
                s2 = xw(h, W2)
      backward  dW2  = hᵀ·g
                dpre = xw_ex(g, W2ᵀ, MASK bits)                ((g·W2ᵀ) ⊙ [h > 0] / (1-p): dropout and ReLU backward)
                dW1  = Xᵀ·dpre,  db1 = Σ_rows dpre,  dX = dpre·W1ᵀ (when X needs it)
    h > 0 holds exactly where the unit was kept by the dropout AND passed the ReLU, so one bit per element,
    written by the first product's epilogue, is the whole backward mask (1/32 of a word per element instead of
    re-reading the activation)."""

    @staticmethod
    def forward(ctx, X, W1, b1, W2, p, seed, training, precision, out):
        p_eff = float(p) if training else 0.0
        bits = torch.empty((X.shape[0], (W1.shape[1] + 31) // 32), dtype=torch.int32, device=X.device)
        h = gemm_raw(X.detach(), W1.detach(), bias=None if b1 is None else b1.detach(), relu=True,
                     precision=precision, dropout=(p_eff, seed), mask_out=bits)
        s2 = gemm_raw(h, W2.detach(), precision=precision, out=out)
        ctx.p, ctx.precision, ctx.has_bias = p_eff, precision, b1 is not None
        ctx.save_for_backward(X, W1, W2, h, bits)
        return s2

    @staticmethod
    def backward(ctx, g):
        X, W1, W2, h, bits = ctx.saved_tensors
        g = g.to(torch.float32)
        if g.stride(1) != 1:
            g = g.contiguous()
        gX = gW1 = gb1 = gW2 = None
        if ctx.needs_input_grad[3]:
            gW2 = _dw(h, g, ctx.precision)
        dpre = gemm_raw(g, W2.detach(), trans_b=True, precision=ctx.precision, mask=(bits, ctx.p))
        if ctx.needs_input_grad[1] and ctx.has_bias and ctx.needs_input_grad[2]:
            gW1, gb1 = dw_colsum(X.detach(), dpre, ctx.precision)    # one pass over dpre for both
        else:
            if ctx.needs_input_grad[1]:
                gW1 = _dw(X.detach(), dpre, ctx.precision)
            if ctx.has_bias and ctx.needs_input_grad[2]:
                gb1 = dpre.sum(dim=0)
        if ctx.needs_input_grad[0]:
            gX = gemm_raw(dpre, W1.detach(), trans_b=True, precision=ctx.precision)
        return gX, gW1, gb1, gW2, None, None, None, None, None


def dw_colsum(X, g, precision=PRECISION_3XTF32):
    """``(Xᵀ·g, g.sum(0))`` from one pass over the two vertex-sized operands (gcnb_xw_tn_colsum): the weight and bias
    gradients of ``X @ W + b``.  Deterministic (fixed summation order)."""
    if not (X.is_cuda and g.is_cuda and X.dtype == torch.float32 and g.dtype == torch.float32 and X.dim() == 2
            and g.dim() == 2 and X.shape[0] == g.shape[0]):
        raise ValueError("gemm.dw_colsum: X [K, M] and g [K, N] must be float32 CUDA matrices over the same rows")
    if X.stride(1) != 1:
        X = X.contiguous()
    if g.stride(1) != 1:
        g = g.contiguous()
    K, M = X.shape
    N = g.shape[1]
    dW = torch.empty(M, N, device=X.device, dtype=torch.float32)
    db = torch.empty(N, device=X.device, dtype=torch.float32)
    stream = ctypes.c_void_p(torch.cuda.current_stream(X.device).cuda_stream)
    with torch.cuda.device(X.device):
        _ffi.call("gcnb_xw_tn_colsum", ctypes.c_void_p(X.data_ptr()), X.stride(0), ctypes.c_void_p(g.data_ptr()),
                  g.stride(0), ctypes.c_void_p(dW.data_ptr()), dW.stride(0), ctypes.c_void_p(db.data_ptr()), M, N, K,
                  int(precision), stream)
    return dW, db


def _dw(X, g, precision):
    """Xᵀ·g with the narrower operand second (see _XW.backward)."""
    if X.shape[1] < g.shape[1] and _SWAP_DW:
        return gemm_raw(g, X, trans_a=True, precision=precision).t().contiguous()
    return gemm_raw(X, g, trans_a=True, precision=precision)


def mlp2(X, W1, b1, W2, p=0.5, seed=0, training=True, precision=PRECISION_3XTF32, out=None):
    """``dropout(relu(X @ W1 + b1), p) @ W2`` with autograd, three fused tensor-core products forward+backward
    plus the two weight-gradient products (see :class:`_MLP2`)."""
    return _MLP2.apply(X, W1, b1, W2, float(p), int(seed), bool(training), int(precision), out)


class _XW(torch.autograd.Function):
    @staticmethod
    def forward(ctx, X, W, bias, relu, precision, out=None):
        if out is not None:
            if relu:
                raise ValueError("gemm.xw: `out` cannot be combined with relu (the result is kept for backward)")
            if (out.dtype != torch.float32 or out.device != X.device or out.dim() != 2 or out.stride(1) != 1
                    or tuple(out.shape) != (X.shape[0], W.shape[1])):
                raise ValueError("gemm.xw: `out` must be a float32 (M, N) tensor on the operands' device with "
                                 "contiguous rows")
            # (no ctx.mark_dirty: `out` is a scratch destination that carries no history — typically a slice of
            # an exchange slot — and marking a view dirty re-bases its autograd history onto the slot's base
            # tensor, which cuts the gradient path to X and W)
        out = gemm_raw(X.detach(), W.detach(), bias=None if bias is None else bias.detach(),
                       relu=relu, precision=precision, out=out)
        ctx.relu = relu
        ctx.precision = precision
        ctx.has_bias = bias is not None
        ctx.save_for_backward(X, W, out if relu else None)
        return out

    @staticmethod
    def backward(ctx, g):
        X, W, out = ctx.saved_tensors
        if ctx.relu:
            g = torch.ops.aten.threshold_backward(g, out, 0.0)   # one pass: g where out > 0 else 0
        g = g.to(torch.float32)
        gX = gW = gb = None
        if ctx.needs_input_grad[0]:
            gX = gemm_raw(g, W.detach(), trans_b=True, precision=ctx.precision)      # dS · Wᵀ
        if ctx.needs_input_grad[1] and ctx.has_bias and ctx.needs_input_grad[2]:
            gW, gb = dw_colsum(X.detach(), g, ctx.precision)
            return gX, gW, gb, None, None, None
        if ctx.needs_input_grad[1]:
            # Xᵀ·dS.  gcnb_xw lays its second operand out once per call (prep kernel, cost ∝ its size), so
            # the narrower of X [M×K] and dS [M×N] goes second: dW = Xᵀ·dS or (dSᵀ·X)ᵀ
            if X.shape[1] < g.shape[1] and _SWAP_DW:
                gW = gemm_raw(g, X.detach(), trans_a=True, precision=ctx.precision).t().contiguous()
            else:
                gW = gemm_raw(X.detach(), g, trans_a=True, precision=ctx.precision)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = g.sum(dim=0)
        return gX, gW, gb, None, None, None


def xw(input: torch.Tensor, weight: torch.Tensor, bias=None, relu=False,
       precision=PRECISION_3XTF32, out=None) -> torch.Tensor:
    """``input @ weight (+ bias)(ReLU)`` with autograd, on tcgen05 tensor cores.  `out`: write the
    product into an existing float32 buffer (e.g. the exchange slot of the row-partitioned SpMM)."""
    return _XW.apply(input, weight, bias, bool(relu), int(precision), out)

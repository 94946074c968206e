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
             precision=PRECISION_3XTF32, out=None) -> torch.Tensor:
    """D = op(A)·op(B) (+bias)(ReLU) on tensor cores, no autograd.  A, B float32 CUDA matrices
    (row-major, any leading dimension); op = transpose when the flag is set."""
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
    stream = ctypes.c_void_p(torch.cuda.current_stream(A.device).cuda_stream)
    with torch.cuda.device(A.device):
        _ffi.call("gcnb_xw", ctypes.c_void_p(A.data_ptr()), A.stride(0), int(trans_a),
                  ctypes.c_void_p(B.data_ptr()), B.stride(0), int(trans_b),
                  ctypes.c_void_p(out.data_ptr()), out.stride(0), M, N, K, bias_ptr, epi,
                  int(precision), stream)
    return out


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

"""SpMM operator of the GCN aggregation path: ``C = Â · B`` forward, ``dB = Âᵀ · dC`` backward.

Replaces the reference's ``flexspmm`` autograd.Function (pygcn/gcn6.py:34-62) and the
``torch.spmm(adj, support)`` call of the torch-API layers (pygcn/gcn5.py:90, gcn1.py:53).  All
arithmetic happens in libgcn_b200.so (hand-written sm_100a kernels, gcn_b200/csrc/spmm*.cu)
through the C ABI of include/gcn_b200.h; torch only owns the device buffers and the stream.
There is no CPU path: operands must live on a CUDA device, anything else raises.

Differences from the reference's Function, on purpose:
  * backward multiplies by Âᵀ (the reference reuses Â, which is only right for a symmetric
    matrix — SURVEY.md §4 defect 3); when the plan finds Â bit-identical to its transpose it runs
    on Â itself, which is the reference's fast path.
  * no per-call ``torch.zeros`` of the output and no ``deepcopy(next_seg)`` (gcn6.py:37-38): the
    kernels write every output element exactly once and hold no mutable launch state.
  * bias add and ReLU (gcn6.py:142, :245) can be fused into the kernel epilogue.
"""
from __future__ import annotations

import ctypes
import weakref

import numpy as np
import torch

from . import _ffi
from ._ffi import np_ptr

_DTYPES = {torch.float32: _ffi.F32, torch.bfloat16: _ffi.BF16}


def _stream_ptr(device) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _require_cuda(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(
            f"gcn_b200: {what} is on '{t.device}'. The SpMM engine is CUDA (sm_100a) only and has "
            "no CPU fallback; move the operands to a B200.")


def _row_major(t: torch.Tensor) -> torch.Tensor:
    """A view/copy of a 2-D tensor whose rows are contiguous (unit column stride)."""
    if t.dim() != 2:
        raise ValueError(f"expected a 2-D dense operand, got shape {tuple(t.shape)}")
    if t.stride(1) != 1 or t.stride(0) < t.shape[1]:
        t = t.contiguous()
    return t


class SpmmPlan:
    """Device-resident sparse matrix prepared for repeated SpMM (gcnb_plan_* in the C ABI).

    Replaces the per-fit tile build + upload of the reference (pygcn/gcn6.py:334-365).  Built from
    a host CSR (int32 indices, fp32 values, ascending columns within a row).
    """

    def __init__(self, rowptr, col, val, shape, device=None, assume_symmetric=False,
                 need_transpose=True, panel=None):
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int32)
        col = np.ascontiguousarray(col, dtype=np.int32)
        val = np.ascontiguousarray(val, dtype=np.float32)
        m, n = int(shape[0]), int(shape[1])
        if len(rowptr) != m + 1:
            raise ValueError("rowptr length differs from shape[0] + 1")
        dev = torch.device("cuda" if device is None else device)
        if dev.type != "cuda":
            raise RuntimeError("gcn_b200: SpmmPlan needs a CUDA device (no CPU fallback)")
        index = dev.index if dev.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", index)
        flags = _ffi.PLAN_DEFAULT
        if assume_symmetric:
            flags |= _ffi.PLAN_ASSUME_SYMMETRIC
        elif not need_transpose:
            flags |= _ffi.PLAN_NO_TRANSPOSE
        if panel is not None:   # None: the library decides (mean degree and measured column sharing)
            flags |= _ffi.PLAN_PANEL if panel else _ffi.PLAN_NO_PANEL
        handle = ctypes.c_void_p()
        torch.cuda.init()
        _ffi.call("gcnb_plan_create", np_ptr(rowptr), np_ptr(col), np_ptr(val), m, n, len(col),
                  flags, index, ctypes.byref(handle))
        self._handle = handle
        self.shape = (m, n)
        self.nnz = len(col)
        self._finalizer = weakref.finalize(self, _ffi.lib.gcnb_plan_destroy, handle)

    # ---- constructors -------------------------------------------------------------------------
    @classmethod
    def from_scipy(cls, mx, device=None, **kw):
        mx = mx.tocsr()
        mx.sort_indices()
        return cls(mx.indptr, mx.indices, mx.data, mx.shape, device=device, **kw)

    @classmethod
    def from_torch(cls, adj: torch.Tensor, device=None, **kw):
        """From a torch sparse COO / CSR tensor (what utils.sparse_mx_to_torch_sparse_tensor and
        adj_norm.to_sparse_csr() produce, pygcn/gcnio/util/utils.py:243-250, gcn6.py:302-312)."""
        if adj.layout == torch.sparse_coo:
            csr = adj.coalesce().to_sparse_csr()
        elif adj.layout == torch.sparse_csr:
            csr = adj
        else:
            raise TypeError(f"adj must be a torch sparse COO or CSR tensor, got layout {adj.layout}")
        if device is None:
            device = adj.device
        return cls(csr.crow_indices().cpu().numpy(), csr.col_indices().cpu().numpy(),
                   csr.values().to(torch.float32).cpu().numpy(), tuple(csr.shape), device=device, **kw)

    # ---- queries ------------------------------------------------------------------------------
    def info(self) -> dict:
        buf = (ctypes.c_int64 * 16)()
        _ffi.call("gcnb_plan_info", self._handle, buf)
        keys = ("m", "n", "nnz", "symmetric", "has_transpose", "rows_warp_class", "rows_block_class",
                "block_items", "max_row_nnz", "device_bytes", "n_sm", "rows_split",
                "stream_segments", "stream_rows_split", "panel_items", "panel_distinct_columns")
        return {k: int(buf[i]) for i, k in enumerate(keys)}

    @property
    def symmetric(self) -> bool:
        return bool(self.info()["symmetric"])

    # ---- raw launches (no autograd) -------------------------------------------------------------
    def _check(self, B: torch.Tensor, rows: int):
        _require_cuda(B, "the dense operand")
        if B.device != self.device:
            raise RuntimeError(f"gcn_b200: operand on {B.device}, plan on {self.device}")
        if B.dtype not in _DTYPES:
            raise TypeError(f"gcn_b200: dense operand must be float32 or bfloat16, got {B.dtype}")
        B = _row_major(B)
        if B.shape[0] != rows:
            raise ValueError(f"dense operand has {B.shape[0]} rows, the sparse matrix needs {rows}")
        return B

    def _out(self, out, rows, k):
        if out is None:
            return torch.empty((rows, k), dtype=torch.float32, device=self.device)
        if out.dtype != torch.float32 or out.shape != (rows, k) or out.stride(1) != 1 or not out.is_cuda:
            raise ValueError("out must be a float32 CUDA tensor of shape (rows, k) with contiguous rows")
        return out

    def spmm(self, B: torch.Tensor, bias: torch.Tensor | None = None, relu: bool = False,
             out: torch.Tensor | None = None, accumulate: bool = False) -> torch.Tensor:
        """C[m×k] = A·B (+ bias) (ReLU); fp32 result.  gcnb_spmm.  accumulate: C = out + A·B."""
        B = self._check(B, self.shape[1])
        k = B.shape[1]
        C = self._out(out, self.shape[0], k)
        epi = _ffi.EPI_NONE
        bias_ptr = None
        if bias is not None:
            if bias.dtype != torch.float32 or bias.numel() != k or bias.device != self.device:
                raise ValueError("bias must be a float32 vector of length k on the plan's device")
            bias = bias.contiguous()
            bias_ptr = ctypes.c_void_p(bias.data_ptr())
            epi |= _ffi.EPI_BIAS
        if relu:
            epi |= _ffi.EPI_RELU
        if accumulate:
            if out is None:
                raise ValueError("accumulate needs an existing `out`")
            epi |= _ffi.EPI_ACCUM
        _ffi.call("gcnb_spmm", self._handle, ctypes.c_void_p(B.data_ptr()), _DTYPES[B.dtype],
                  B.stride(0), ctypes.c_void_p(C.data_ptr()), C.stride(0), k, bias_ptr, epi,
                  _stream_ptr(self.device))
        return C

    def spmm_t(self, B: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
        """C[n×k] = Aᵀ·B; fp32 result.  gcnb_spmm_t."""
        B = self._check(B, self.shape[0])
        k = B.shape[1]
        C = self._out(out, self.shape[1], k)
        _ffi.call("gcnb_spmm_t", self._handle, ctypes.c_void_p(B.data_ptr()), _DTYPES[B.dtype],
                  B.stride(0), ctypes.c_void_p(C.data_ptr()), C.stride(0), k,
                  _stream_ptr(self.device))
        return C

    def spmm_host(self, B_host: np.ndarray, transpose: bool = False, out: np.ndarray | None = None):
        """End-to-end call on HOST buffers (H2D, SpMM, D2H inside): gcnb_spmm_host."""
        B_host = np.ascontiguousarray(B_host, dtype=np.float32)
        rows_in, rows_out = (self.shape[0], self.shape[1]) if transpose else (self.shape[1], self.shape[0])
        if B_host.ndim != 2 or B_host.shape[0] != rows_in:
            raise ValueError(f"host operand must have shape ({rows_in}, k)")
        k = B_host.shape[1]
        if out is None:
            out = np.empty((rows_out, k), dtype=np.float32)
        elif out.dtype != np.float32 or out.shape != (rows_out, k) or not out.flags.c_contiguous:
            raise ValueError("out must be a C-contiguous float32 array of shape (rows, k)")
        with torch.cuda.device(self.device):
            _ffi.call("gcnb_spmm_host", self._handle, np_ptr(B_host), np_ptr(out), k, int(transpose),
                      _stream_ptr(self.device))
        return out


def permute_rows(X: torch.Tensor, vomp: torch.Tensor) -> torch.Tensor:
    """out[r,:] = X[vomp[r],:] — the net effect of the reference's permutate() (permutate.cu:40-59),
    out of place.  X float32 [n×k], vomp int32 [n] (``vomp[new] = old``), both on the GPU."""
    _require_cuda(X, "X")
    _require_cuda(vomp, "vomp")
    if X.dtype != torch.float32 or vomp.dtype != torch.int32:
        raise TypeError("permute_rows needs float32 X and int32 vomp")
    X = X.contiguous()
    vomp = vomp.contiguous()
    if X.dim() != 2 or vomp.numel() != X.shape[0]:
        raise ValueError("vomp must have one entry per row of X")
    out = torch.empty_like(X)
    _ffi.call("gcnb_permute_rows", ctypes.c_void_p(X.data_ptr()), ctypes.c_void_p(vomp.data_ptr()),
              ctypes.c_void_p(out.data_ptr()), X.shape[0], X.shape[1], _stream_ptr(X.device))
    return out


# ---- plan cache keyed on the adjacency tensor -----------------------------------------------------
_PLAN_ATTR = "_gcnb_plan"


def plan_for(adj, device=None, **kw) -> SpmmPlan:
    """The SpmmPlan of ``adj``: a plan passes through; a torch sparse tensor gets one built on
    first use and cached on the tensor object (the layer is called every epoch with the same
    ``adj_norm``, pygcn/gcn5.py:226-262)."""
    if isinstance(adj, SpmmPlan):
        return adj
    if isinstance(adj, torch.Tensor):
        plan = getattr(adj, _PLAN_ATTR, None)
        if plan is None:
            if not adj.is_sparse and adj.layout != torch.sparse_csr:
                raise TypeError("gcn_b200: adj must be a torch sparse tensor or an SpmmPlan "
                                "(a dense adjacency is not an SpMM)")
            _require_cuda(adj, "adj")
            plan = SpmmPlan.from_torch(adj, device=device, **kw)
            setattr(adj, _PLAN_ATTR, plan)
        return plan
    try:
        import scipy.sparse as sp
        if sp.issparse(adj):
            return SpmmPlan.from_scipy(adj, device=device, **kw)
    except ImportError:  # pragma: no cover
        pass
    raise TypeError(f"gcn_b200: cannot build an SpMM plan from {type(adj)}")


class _SpMM(torch.autograd.Function):
    """out = A·B (+bias)(ReLU);  dB = Aᵀ·dOut,  dbias = Σ_rows dOut."""

    @staticmethod
    def forward(ctx, plan: SpmmPlan, B, bias, relu: bool):
        out = plan.spmm(B.detach(), None if bias is None else bias.detach(), relu)
        ctx.plan = plan
        ctx.relu = relu
        ctx.b_dtype = B.dtype
        ctx.has_bias = bias is not None
        if relu:
            ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        g = grad_out
        if ctx.relu:
            (out,) = ctx.saved_tensors
            g = torch.ops.aten.threshold_backward(g, out, 0.0)   # one pass: g where out > 0 else 0
        g = g.to(torch.float32)
        grad_B = grad_bias = None
        if ctx.needs_input_grad[1]:
            grad_B = ctx.plan.spmm_t(g).to(ctx.b_dtype)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            grad_bias = g.sum(dim=0)
        return None, grad_B, grad_bias, None


def spmm(adj, B: torch.Tensor, bias: torch.Tensor | None = None, relu: bool = False) -> torch.Tensor:
    """Drop-in for ``torch.spmm(adj, B)`` (pygcn/gcn5.py:90) with autograd through B (and bias).

    ``adj``: torch sparse COO/CSR tensor on the GPU, or an :class:`SpmmPlan`."""
    plan = plan_for(adj)
    _require_cuda(B, "the dense operand")
    return _SpMM.apply(plan, B, bias, bool(relu))


# ---- fused head: log-softmax + NLL on the training vertices (gcnb_logsoftmax_nll_*) -------------------------
class _LogSoftmaxNLL(torch.autograd.Function):
    """(logp, loss) = (log_softmax(logits[:, :ncls]), nll_loss(logp[idx_train], labels[idx_train])) in one pass,
    backward = one kernel writing dlogits in the (possibly padded) layout of `logits`.  Replaces F.log_softmax
    (pygcn/gcn6.py:248) + F.nll_loss (gcn6.py:427) and their autograd chain.  `logp` is returned for the
    caller (the model's output) and carries no gradient."""

    @staticmethod
    def forward(ctx, logits, ncls, labels, idx_train, is_train, denom):
        n = logits.shape[0]
        logp = torch.empty((n, ncls), dtype=torch.float32, device=logits.device)
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        with torch.cuda.device(logits.device):
            _ffi.call("gcnb_logsoftmax_nll_fwd", ctypes.c_void_p(logits.data_ptr()), logits.stride(0), n, int(ncls),
                      ctypes.c_void_p(labels.data_ptr()), ctypes.c_void_p(idx_train.data_ptr()), idx_train.numel(),
                      int(denom), ctypes.c_void_p(logp.data_ptr()), logp.stride(0), ctypes.c_void_p(loss.data_ptr()),
                      _stream_ptr(logits.device))
        ctx.save_for_backward(logp, labels, is_train)
        ctx.ncls, ctx.ncols, ctx.denom = int(ncls), logits.shape[1], int(denom)
        ctx.mark_non_differentiable(logp)
        return logp, loss

    @staticmethod
    def backward(ctx, _g_logp, g_loss):
        logp, labels, is_train = ctx.saved_tensors
        n = logp.shape[0]
        dlogits = torch.empty((n, ctx.ncols), dtype=torch.float32, device=logp.device)
        g = g_loss.to(torch.float32).contiguous()
        with torch.cuda.device(logp.device):
            _ffi.call("gcnb_logsoftmax_nll_bwd", ctypes.c_void_p(logp.data_ptr()), logp.stride(0),
                      ctypes.c_void_p(labels.data_ptr()), ctypes.c_void_p(is_train.data_ptr()),
                      ctypes.c_void_p(g.data_ptr()), ctx.denom, n, ctx.ncls, ctx.ncols,
                      ctypes.c_void_p(dlogits.data_ptr()), dlogits.stride(0), _stream_ptr(logp.device))
        return dlogits, None, None, None, None, None


def logsoftmax_nll(logits: torch.Tensor, ncls: int, labels: torch.Tensor, idx_train: torch.Tensor,
                   is_train: torch.Tensor | None = None, denom: int | None = None):
    """``(F.log_softmax(logits[:, :ncls], 1), F.nll_loss(logp[idx_train], labels[idx_train]))`` fused.
    logits: float32 CUDA [n, >= ncls] with contiguous rows (columns past ncls are padding); labels int64 [n];
    idx_train int64 (distinct vertices); is_train: uint8/bool [n] membership of idx_train (built when omitted);
    denom: divisor of the loss sum (default len(idx_train); the global count on a row partition)."""
    _require_cuda(logits, "logits")
    if logits.dtype != torch.float32 or logits.dim() != 2 or logits.stride(1) != 1 or logits.shape[1] < ncls:
        raise ValueError("logits must be float32 [n, >= ncls] with contiguous rows")
    labels = labels.to(torch.long).contiguous()
    idx_train = idx_train.to(torch.long).contiguous()
    if is_train is None:
        is_train = torch.zeros(logits.shape[0], dtype=torch.uint8, device=logits.device)
        is_train[idx_train] = 1
    elif is_train.dtype == torch.bool:
        is_train = is_train.view(torch.uint8)
    return _LogSoftmaxNLL.apply(logits, int(ncls), labels, idx_train, is_train.contiguous(),
                                idx_train.numel() if denom is None else int(denom))

"""1-D row partition of Â across the GPUs of one box, with an all-gather of the dense operand per
SpMM.  New relative to the reference, which is single-GPU (no NCCL/MPI anywhere).

One process per GPU (``torch.distributed``, NCCL over NVLink 5 / NVSwitch).  Rank p owns the
contiguous row block ``[bounds[p], bounds[p+1])`` of Â (all columns), of the features, activations
and gradients; weights are replicated.  Blocks are balanced by non-zeros, not by rows.

Per SpMM:  the local operand block ``B_p`` is all-gathered on a side stream while the part of
``Â[rows_p, :]`` whose columns are local (``Â_pp``, no communication needed) is multiplied on the
compute stream; when the gather has landed, the remote-column part accumulates into the same
output (``C += Â_p,rest · B``, epilogue flag GCNB_EPI_ACCUM).  With a locality-preserving vertex
order most non-zeros sit in ``Â_pp``, so the gather hides behind it.  Backward is the same
operation on the transposed blocks (``dB_p = Σ_q Â_qpᵀ dC_q`` = row block p of Âᵀ times the
gathered dC), so it also costs one all-gather and no reduce-scatter; for a symmetric Â the
transposed blocks are the forward blocks.

Column ids of the remote part are remapped once, on the host, into the padded layout
``q * stride + (col - bounds[q])`` that ``all_gather_into_tensor`` produces (equal-sized slots of
``stride = max block rows``), so no compaction copy is needed after the gather.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from . import ops


def nnz_balanced_bounds(rowptr, parts: int) -> np.ndarray:
    """Row bounds [parts+1] of contiguous blocks with ≈ nnz/parts non-zeros each (prefix-sum split).
    Block p is rows bounds[p]..bounds[p+1]-1.  Deterministic, identical on every rank."""
    rowptr = np.asarray(rowptr, dtype=np.int64)
    m = len(rowptr) - 1
    nnz = int(rowptr[-1])
    if parts < 1:
        raise ValueError("parts must be >= 1")
    # weight rows by nnz + 1 so that empty rows are spread too
    w = rowptr + np.arange(m + 1, dtype=np.int64)
    targets = (np.arange(1, parts, dtype=np.int64) * int(w[-1])) // parts
    cuts = np.searchsorted(w, targets, side="left")
    bounds = np.concatenate(([0], cuts, [m])).astype(np.int64)
    bounds = np.maximum.accumulate(bounds)
    assert bounds[0] == 0 and bounds[-1] == m and nnz >= 0
    return bounds


def split_local_remote(rowptr, col, val, bounds, rank: int):
    """Row block `rank` of a CSR, split by column ownership.

    Returns (local, remote, stride): `local` = CSR [m_p × m_p] of the columns inside the block
    (re-based to 0); `remote` = CSR [m_p × parts*stride] of all other columns with ids remapped
    to the padded all-gather layout; stride = max block rows."""
    rowptr = np.asarray(rowptr, dtype=np.int64)
    col = np.asarray(col)
    val = np.asarray(val, dtype=np.float32)
    parts = len(bounds) - 1
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    stride = int(np.max(np.diff(bounds))) if parts > 0 else 0
    z0, z1 = int(rowptr[lo]), int(rowptr[hi])
    c = col[z0:z1].astype(np.int64)
    v = val[z0:z1]
    rows = np.repeat(np.arange(hi - lo, dtype=np.int64), np.diff(rowptr[lo:hi + 1]))
    is_local = (c >= lo) & (c < hi)

    def build(mask, newcol):
        counts = np.bincount(rows[mask], minlength=hi - lo)
        rp = np.zeros(hi - lo + 1, dtype=np.int32)
        np.cumsum(counts, out=rp[1:])
        return rp, newcol.astype(np.int32), v[mask].copy()

    local = build(is_local, c[is_local] - lo)
    cr = c[~is_local]
    owner = np.searchsorted(bounds, cr, side="right") - 1
    remote = build(~is_local, owner * stride + (cr - bounds[owner]))
    return local, remote, stride


class PartitionedSpMM:
    """Row block `rank` of Â (and of Âᵀ unless symmetric) prepared for the gathered SpMM."""

    def __init__(self, rowptr, col, val, n: int, group=None, device=None, bounds=None,
                 symmetric: bool | None = None):
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.n = int(n)
        rowptr = np.asarray(rowptr)
        if len(rowptr) != self.n + 1:
            raise ValueError("row-partitioned SpMM needs a square matrix (rows = vertices)")
        self.bounds = nnz_balanced_bounds(rowptr, self.world) if bounds is None else np.asarray(bounds)
        self.lo, self.hi = int(self.bounds[self.rank]), int(self.bounds[self.rank + 1])
        self.m_local = self.hi - self.lo
        self.device = torch.device("cpu" if device is None else device)
        if symmetric is None:
            symmetric = _is_symmetric(rowptr, col, val, self.n)
        self.symmetric = bool(symmetric)
        self.fwd = self._make_blocks(rowptr, col, val)
        if self.symmetric:
            self.bwd = self.fwd
        else:
            t_rowptr, t_col, t_val = _transpose(rowptr, col, val, self.n)
            self.bwd = self._make_blocks(t_rowptr, t_col, t_val)
        self._side = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None

    def _make_blocks(self, rowptr, col, val):
        local, remote, stride = split_local_remote(rowptr, col, val, self.bounds, self.rank)
        self.stride = stride
        blocks = {"local_csr": local, "remote_csr": remote}
        if self.device.type == "cuda":
            # forward-only plans: the transposed blocks are separate plans (self.bwd)
            blocks["local"] = ops.SpmmPlan(*local, (self.m_local, self.m_local), device=self.device,
                                           need_transpose=False)
            blocks["remote"] = ops.SpmmPlan(*remote, (self.m_local, self.world * stride),
                                            device=self.device, need_transpose=False)
        return blocks

    # the two kernels; tests on CPU (gloo) substitute these to exercise the exchange logic
    def _mul_local(self, blocks, B_local, out):
        blocks["local"].spmm(B_local, out=out)

    def _mul_remote_accumulate(self, blocks, B_gathered, out):
        blocks["remote"].spmm(B_gathered, out=out, accumulate=True)

    def _gathered(self, blocks, B_local):
        k = B_local.shape[1]
        padded = B_local
        if self.m_local != self.stride:
            padded = B_local.new_zeros((self.stride, k))
            padded[: self.m_local] = B_local
        full = B_local.new_empty((self.world * self.stride, k))
        dist.all_gather_into_tensor(full, padded.contiguous(), group=self.group)
        return full

    def _apply(self, blocks, B_local: torch.Tensor) -> torch.Tensor:
        if B_local.shape[0] != self.m_local:
            raise ValueError(f"rank {self.rank} owns {self.m_local} rows, operand has {B_local.shape[0]}")
        B_local = B_local.contiguous()
        out = torch.empty((self.m_local, B_local.shape[1]), dtype=torch.float32, device=B_local.device)
        if self._side is not None:
            cur = torch.cuda.current_stream(self.device)
            self._side.wait_stream(cur)
            with torch.cuda.stream(self._side):
                full = self._gathered(blocks, B_local)
            self._mul_local(blocks, B_local, out)          # overlaps the all-gather
            cur.wait_stream(self._side)
            full.record_stream(cur)
        else:
            full = self._gathered(blocks, B_local)
            self._mul_local(blocks, B_local, out)
        self._mul_remote_accumulate(blocks, full, out)
        return out

    def forward(self, B_local):
        """Rows [lo, hi) of Â·B, from the local rows of B."""
        return self._apply(self.fwd, B_local)

    def backward(self, G_local):
        """Rows [lo, hi) of Âᵀ·G, from the local rows of G."""
        return self._apply(self.bwd, G_local)


class _PartSpMM(torch.autograd.Function):
    @staticmethod
    def forward(ctx, part: PartitionedSpMM, B_local):
        ctx.part = part
        return part.forward(B_local.detach())

    @staticmethod
    def backward(ctx, g):
        return None, ctx.part.backward(g.contiguous())


def spmm(part: PartitionedSpMM, B_local: torch.Tensor) -> torch.Tensor:
    """Row-partitioned ``Â·B`` with autograd (each rank passes and receives its own row block)."""
    return _PartSpMM.apply(part, B_local)


def allreduce_grads(module: torch.nn.Module, group=None) -> None:
    """Sum replicated-weight gradients over ranks (the [in×out] all-reduce of SURVEY.md §8e)."""
    for p in module.parameters():
        if p.grad is not None:
            dist.all_reduce(p.grad, op=dist.ReduceOp.SUM, group=group)


def _transpose(rowptr, col, val, n):
    import scipy.sparse as sp
    a = sp.csr_matrix((np.asarray(val, dtype=np.float32), np.asarray(col), np.asarray(rowptr)), shape=(n, n))
    t = a.T.tocsr()
    t.sort_indices()
    return t.indptr, t.indices, t.data


def _is_symmetric(rowptr, col, val, n) -> bool:
    t_rowptr, t_col, t_val = _transpose(rowptr, col, val, n)
    return (np.array_equal(t_rowptr, rowptr) and np.array_equal(t_col, col)
            and np.array_equal(t_val, np.asarray(val, dtype=np.float32)))


class PartitionedGCN(torch.nn.Module):
    """The reference's 2-layer GCN (pygcn/gcn5.py:146-196: gc1 = A·(X·W1)+b1, ReLU, dropout,
    gc2 = (A·H)·W2+b2, log_softmax) on the row partition: every rank holds its row block of the
    features, activations and labels; W1/W2/biases are replicated and their gradients all-reduced
    (SURVEY.md §8e).  One epoch = 4 gathered SpMMs (2 forward, 2 backward) + 6 dense products."""

    def __init__(self, part: PartitionedSpMM, nfeat, nhid, nclass, dropout=0.5, lr=0.01,
                 weight_decay=5e-4, seed=15):
        super().__init__()
        from .layers import GraphConvolution, GraphConvolution2
        self.part = part
        g = torch.Generator().manual_seed(seed)   # identical replicas on every rank
        self.gc1 = GraphConvolution(nfeat, nhid)
        self.gc2 = GraphConvolution2(nhid, nclass)
        for p in self.parameters():
            stdv = 1.0 / (p.shape[-1] ** 0.5)
            p.data.copy_((torch.rand(p.shape, generator=g) * 2 - 1) * stdv)
        self.dropout = dropout
        self.to(part.device)
        self.optimizer = torch.optim.Adam(self.parameters(), lr=lr, weight_decay=weight_decay)

    def forward(self, x_local):
        from . import gemm
        support = gemm.xw(x_local, self.gc1.weight)
        h = torch.relu(spmm(self.part, support) + self.gc1.bias)
        h = torch.nn.functional.dropout(h, self.dropout, training=self.training)
        s2 = spmm(self.part, h)
        out = gemm.xw(s2, self.gc2.weight, bias=self.gc2.bias)
        return torch.nn.functional.log_softmax(out, dim=1)

    def step(self, x_local, labels_local, idx_train_local, n_train_global):
        """One training epoch (forward, NLL on the local training rows, backward, gradient
        all-reduce, Adam).  Returns the local loss contribution (sum / global count)."""
        self.train()
        self.optimizer.zero_grad(set_to_none=True)
        out = self.forward(x_local)
        loss = torch.nn.functional.nll_loss(out[idx_train_local], labels_local[idx_train_local],
                                            reduction="sum") / n_train_global
        loss.backward()
        allreduce_grads(self, self.part.group)
        self.optimizer.step()
        return loss

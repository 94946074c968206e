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

Two ways of bringing the remote rows in (``exchange=``):

``"allgather"``  NCCL ``all_gather_into_tensor`` of the padded blocks.  Column ids of the remote part
    are remapped once, on the host, into the padded layout ``q * stride + (col - bounds[q])`` that
    the collective produces (equal-sized slots of ``stride = max block rows``), so no compaction
    copy is needed after the gather.
``"pull"``  our own kernel over NVLink peer memory (``gcnb_pull_rows``, gcn_b200/csrc/peer.cu).  Every
    rank publishes its block in an exchange buffer the peers have mapped through CUDA IPC; the
    *halo* — the distinct remote rows this rank's non-zeros touch, listed once at plan time and
    sorted by (owner, row) — is read straight out of the owners' HBM into a compact buffer, and the
    remote part's column ids index that buffer.  Only rows that are needed cross NVLink, once.
    Ordering between ranks: a one-element all-reduce on the exchange stream after the block is
    published (so every peer's block is complete before anybody pulls); the exchange buffer is
    double-buffered, which makes that same all-reduce of the NEXT exchange the proof that all
    pulls of this one have finished before its slot is overwritten.
"""
from __future__ import annotations

import os

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


def compact_halo(remote, bounds, stride: int):
    """Re-index the remote part of a row block by its halo.

    `remote` is the CSR returned by :func:`split_local_remote` (column ids in the padded layout
    ``owner * stride + row_in_owner``).  Returns (csr, halo_part, halo_row): the same CSR with
    column ids replaced by positions in the sorted list of DISTINCT remote rows, and that list as
    (owner rank, row inside the owner's block) int32 arrays — sorted by owner, then row, which is
    also ascending global id, so columns stay ascending within a row."""
    rp, col, val = remote
    uniq, inv = np.unique(col.astype(np.int64), return_inverse=True)
    halo_part = (uniq // max(stride, 1)).astype(np.int32)
    halo_row = (uniq - halo_part.astype(np.int64) * stride).astype(np.int32)
    return (rp, inv.astype(np.int32), val), halo_part, halo_row


class _PeerExchange:
    """Double-buffered exchange buffer of one rank for one (row width, dtype), mapped on every peer.

    Collective to construct (handles travel by ``all_gather_object``)."""

    def __init__(self, rows: int, k: int, device, group, dtype=torch.float32):
        import ctypes
        from . import _ffi
        self._ffi, self._ct = _ffi, ctypes
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.device, self.rows, self.k = device, int(rows), int(k)
        self.dtype = dtype
        self.esize = torch.empty(0, dtype=dtype).element_size()
        self.slot_bytes = ((self.rows * self.k * self.esize + 255) // 256) * 256
        ptr = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        _ffi.call("gcnb_peer_alloc", device.index, 2 * self.slot_bytes, ctypes.byref(ptr), handle)
        self.own_ptr = ptr.value
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=group)
        self.base = []          # device pointer of every rank's buffer in THIS process
        self._opened = []
        for q, h in enumerate(handles):
            if q == self.rank:
                self.base.append(self.own_ptr)
                continue
            p = ctypes.c_void_p()
            _ffi.call("gcnb_peer_open", device.index, (ctypes.c_ubyte * 64).from_buffer_copy(h), ctypes.byref(p))
            self.base.append(p.value)
            self._opened.append(p.value)
        self.own = [self._wrap(self.own_ptr + s * self.slot_bytes) for s in (0, 1)]
        self.turn = 0

    def _wrap(self, ptr):
        class _Raw:   # CUDA array interface view of the raw allocation (the library owns the memory)
            pass
        raw = _Raw()
        # bfloat16 has no type string in the array interface: map the bytes as int16 and reinterpret
        typestr = {4: "<f4", 2: "<i2"}[self.esize]
        raw.__cuda_array_interface__ = {"shape": (self.rows, self.k), "typestr": typestr, "data": (ptr, False),
                                        "version": 2, "strides": None}
        t = torch.as_tensor(raw, device=self.device)
        if t.dtype != self.dtype:
            t = t.view(self.dtype)
        assert t.data_ptr() == ptr
        return t

    def bases(self, slot: int):
        arr = (self._ct.c_void_p * self.world)(*[b + slot * self.slot_bytes for b in self.base])
        return arr

    def close(self):
        for p in self._opened:
            self._ffi.lib.gcnb_peer_close(self.device.index, self._ct.c_void_p(p))
        self._opened = []
        if self.own_ptr:
            self._ffi.lib.gcnb_peer_free(self.device.index, self._ct.c_void_p(self.own_ptr))
            self.own_ptr = None


_LOCAL_AFTER_BARRIER = os.environ.get("GCNB_LOCAL_AFTER_BARRIER", "auto")   # "0", "1" or "auto"


class _PeerFlags:
    """One array of `world` uint32 words per rank in peer-mapped memory, for gcnb_peer_barrier.  Collective to
    construct: the words are zeroed BEFORE the handles travel, so no peer can write into an array that its owner
    has yet to clear."""

    def __init__(self, device, group):
        import ctypes
        from . import _ffi
        self._ffi, self._ct = _ffi, ctypes
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.device = device
        ptr = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        _ffi.call("gcnb_peer_alloc", device.index, 256, ctypes.byref(ptr), handle)
        self.own_ptr = ptr.value

        class _Raw:
            pass
        raw = _Raw()
        raw.__cuda_array_interface__ = {"shape": (64,), "typestr": "<i4", "data": (self.own_ptr, False),
                                        "version": 2, "strides": None}
        torch.as_tensor(raw, device=device).zero_()
        torch.cuda.synchronize(device)
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=group)
        self.base, self._opened = [], []
        for q, h in enumerate(handles):
            if q == self.rank:
                self.base.append(self.own_ptr)
                continue
            p = ctypes.c_void_p()
            _ffi.call("gcnb_peer_open", device.index, (ctypes.c_ubyte * 64).from_buffer_copy(h), ctypes.byref(p))
            self.base.append(p.value)
            self._opened.append(p.value)
        self.bases = (ctypes.c_void_p * self.world)(*self.base)
        self.epoch = 0

    def barrier(self, stream):
        self.epoch = (self.epoch + 1) & 0xFFFFFFFF
        self._ffi.call("gcnb_peer_barrier", self.bases, self.world, self.rank, self._ct.c_uint32(self.epoch),
                       self._ct.c_void_p(stream.cuda_stream))

    def close(self):
        for p in self._opened:
            self._ffi.lib.gcnb_peer_close(self.device.index, self._ct.c_void_p(p))
        self._opened = []
        if self.own_ptr:
            self._ffi.lib.gcnb_peer_free(self.device.index, self._ct.c_void_p(self.own_ptr))
            self.own_ptr = None


class PartitionedSpMM:
    """Row block `rank` of Â (and of Âᵀ unless symmetric) prepared for the gathered SpMM."""

    def __init__(self, rowptr, col, val, n: int, group=None, device=None, bounds=None,
                 symmetric: bool | None = None, exchange: str = "allgather", wire_dtype=None):
        if exchange not in ("allgather", "pull"):
            raise ValueError("exchange must be 'allgather' or 'pull'")
        if wire_dtype not in (None, torch.float32, torch.bfloat16):
            raise ValueError("wire_dtype must be None / torch.float32 / torch.bfloat16")
        self.exchange = exchange
        # "bf16 on the wire" (SURVEY.md §8e): the rows a rank publishes for its peers are rounded to
        # bfloat16, which halves the NVLink bytes and the gather bytes of the remote-column block; the
        # local-column block still multiplies the rank's own fp32 rows.  The result is then within the
        # bf16-feature tolerance of the north star (1e-2), not the fp32 one, so it is opt-in.
        self.wire_dtype = torch.float32 if wire_dtype is None else wire_dtype
        self._exchanges = {}     # k -> _PeerExchange
        self._halo_bufs = {}     # (id(blocks), k) -> compact halo rows
        self._flag = None
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
        # exchange stream: high priority, so that its (few, short) CTAs are scheduled between the
        # CTAs of the local-block SpMM that fills the machine at the same time
        self._side = torch.cuda.Stream(self.device, priority=-1) if self.device.type == "cuda" else None
        # rendezvous of the ranks: flags in peer memory (one small kernel, gcnb_peer_barrier) when the ranks sit on
        # different GPUs of one box (NCCL group); a one-element all-reduce otherwise (gloo: CPU tests, and the GPU
        # tests that put two ranks on one device, where a spinning kernel would wait for a time slice)
        self._after_barrier = None
        self._peer_flags = None
        if (self.exchange == "pull" and self.device.type == "cuda" and self.world > 1
                and dist.get_backend(group) == "nccl" and os.environ.get("GCNB_FLAG_BARRIER", "1") != "0"):
            self._peer_flags = _PeerFlags(self.device, group)

    def _make_blocks(self, rowptr, col, val):
        local, remote, stride = split_local_remote(rowptr, col, val, self.bounds, self.rank)
        self.stride = stride
        blocks = {"local_csr": local, "remote_csr": remote}
        n_remote_cols = self.world * stride
        if self.exchange == "pull":
            remote, halo_part, halo_row = compact_halo(remote, self.bounds, stride)
            blocks["remote_csr"] = remote
            blocks["halo_part_np"], blocks["halo_row_np"] = halo_part, halo_row
            # every rank starts its pull at the rows of the NEXT rank and wraps around, so that the
            # ranks read from different owners at any moment (see gcnb_pull_rows)
            first = int(np.searchsorted(halo_part, self.rank + 1, side="left"))
            blocks["halo_first"] = first if first < len(halo_part) else 0
            blocks["halo_part"] = torch.from_numpy(halo_part).to(self.device)
            blocks["halo_row"] = torch.from_numpy(halo_row).to(self.device)
            n_remote_cols = max(len(halo_row), 1)
        blocks["n_remote_cols"] = n_remote_cols
        if self.device.type == "cuda":
            # forward-only plans: the transposed blocks are separate plans (self.bwd)
            blocks["local"] = ops.SpmmPlan(*local, (self.m_local, self.m_local), device=self.device,
                                           need_transpose=False)
            blocks["remote"] = ops.SpmmPlan(*remote, (self.m_local, n_remote_cols),
                                            device=self.device, need_transpose=False)
        return blocks

    # the two kernels; tests on CPU (gloo) substitute these to exercise the exchange logic
    def _mul_local(self, blocks, B_local, out):
        blocks["local"].spmm(B_local, out=out)

    def _mul_remote_accumulate(self, blocks, B_gathered, out):
        blocks["remote"].spmm(B_gathered, out=out, accumulate=True)

    def _gathered(self, blocks, B_local):
        k = B_local.shape[1]
        padded = B_local.to(self.wire_dtype)
        if self.m_local != self.stride:
            padded = padded.new_zeros((self.stride, k))
            padded[: self.m_local] = B_local
        full = padded.new_empty((self.world * self.stride, k))
        dist.all_gather_into_tensor(full, padded.contiguous(), group=self.group)
        return full

    # ---- "pull" exchange ------------------------------------------------------------------------
    def _exchange_for(self, k: int):
        ex = self._exchanges.get(k)
        if ex is None:   # collective: every rank reaches its first SpMM of width k together
            ex = self._exchanges[k] = _PeerExchange(self.stride, k, self.device, self.group, self.wire_dtype)
        return ex

    def next_operand(self, k: int):
        """Pull mode: a [m_local × k] float32 view of the free exchange slot for the producer of the
        next operand (``gemm.xw(..., out=part.next_operand(k))``): an operand that already lives there
        is published without a copy.  Collective on first use of a width.  None in all-gather mode."""
        if self.exchange != "pull" or self.wire_dtype != torch.float32:
            return None     # a bf16 slot cannot hold the producer's fp32 result: it is cast on publish
        ex = self._exchange_for(k)
        slot = ex.turn
        ex.turn ^= 1
        return ex.own[slot][: self.m_local]

    def _publish(self, ex, B_local):
        """Make the local block visible to the peers; returns the slot it is in.  A block that was
        produced in place (next_operand) is already there; anything else is copied into the free slot."""
        for slot in (0, 1):
            if (B_local.dtype == ex.dtype and B_local.data_ptr() == ex.own[slot].data_ptr()
                    and B_local.stride(0) == ex.k):
                return slot
        slot = ex.turn
        ex.turn ^= 1
        ex.own[slot][: self.m_local].copy_(B_local)     # rounds to the wire dtype when they differ
        return slot

    def _rank_barrier(self):
        """Stream-ordered rendezvous of all ranks on the current stream: flags in peer memory
        (gcnb_peer_barrier) between the GPUs of one box, else a one-element all-reduce."""
        if self._peer_flags is not None:
            self._peer_flags.barrier(torch.cuda.current_stream(self.device))
            return
        if self._flag is None:
            self._flag = torch.zeros(1, dtype=torch.float32, device=self.device)
        dist.all_reduce(self._flag, group=self.group)

    def _pull(self, blocks, ex, slot, k):
        """The halo rows of `blocks`, compact, read out of the owners' exchange buffers."""
        import ctypes
        from . import _ffi
        n_halo = blocks["halo_row"].numel()
        key = (id(blocks), k)
        halo = self._halo_bufs.get(key)
        if halo is None:
            halo = self._halo_bufs[key] = torch.empty((blocks["n_remote_cols"], k), dtype=self.wire_dtype,
                                                      device=self.device)
        if n_halo:
            row_bytes = k * halo.element_size()
            _ffi.call("gcnb_pull_rows", ex.bases(slot), self.world, row_bytes,
                      ctypes.c_void_p(blocks["halo_part"].data_ptr()),
                      ctypes.c_void_p(blocks["halo_row"].data_ptr()), ctypes.c_void_p(halo.data_ptr()),
                      row_bytes, n_halo, row_bytes, blocks["halo_first"],
                      ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
        return halo

    def _exchange(self, blocks, B_local, pre=None):
        """Remote rows of the operand in the layout the remote block's columns index.  `pre` is what
        _prepare returned on the compute stream (the published slot in pull mode)."""
        if self.exchange == "allgather":
            return self._gathered(blocks, B_local)
        ex, slot = pre
        self._rank_barrier()
        if self.device.type == "cuda":
            if self._after_barrier is None:
                self._after_barrier = torch.cuda.Event()
            self._after_barrier.record(torch.cuda.current_stream(self.device))
        return self._pull(blocks, ex, slot, B_local.shape[1])

    def _local_after_barrier(self, blocks, B_local) -> bool:
        """Should the local block wait for the rendezvous kernel?  Launched first, it fills every SM and the
        rendezvous takes 0.1-0.2 ms to get through instead of 0.01-0.02 (profiles/r02_ar_timeline_n2.jsonl) — which
        matters when rendezvous + pull is the critical path (8 GPUs) and costs when the local block is (2 GPUs: it
        runs 0.4 ms longer beside a pull that started with it).  'auto' compares rough estimates of the two: halo
        bytes at 550 GB/s against local non-zeros at (5 + k/4) ps each, both measured on the products shape, and
        waits when the pull is 1.3x longer (8 GPUs, profiles/r02_as_*: k = 100, ratio 1.4: step 1.14 -> 1.07 ms;
        the epoch's k = 48 products, ratio 1.2: 2.85 -> 3.00 ms per epoch, so those do not wait)."""
        if _LOCAL_AFTER_BARRIER in ("0", "1"):
            return _LOCAL_AFTER_BARRIER == "1"
        if self.exchange != "pull":
            return False
        k = B_local.shape[1]
        esize = 2 if self.wire_dtype == torch.bfloat16 else 4
        pull_s = blocks["halo_row"].numel() * k * esize / 550e9
        local_s = int(blocks["local_csr"][0][-1]) * (5.0 + 0.25 * k) * 1e-12
        return pull_s > 1.3 * local_s

    def _prepare(self, B_local):
        if self.exchange == "allgather":
            return None
        ex = self._exchange_for(B_local.shape[1])
        return ex, self._publish(ex, B_local)

    def close(self):
        for ex in self._exchanges.values():
            ex.close()
        self._exchanges = {}
        if self._peer_flags is not None:
            self._peer_flags.close()
            self._peer_flags = None

    def _apply(self, blocks, B_local: torch.Tensor) -> torch.Tensor:
        if B_local.shape[0] != self.m_local:
            raise ValueError(f"rank {self.rank} owns {self.m_local} rows, operand has {B_local.shape[0]}")
        B_local = B_local.contiguous()
        out = torch.empty((self.m_local, B_local.shape[1]), dtype=torch.float32, device=B_local.device)
        pre = self._prepare(B_local)
        if self._side is not None:
            cur = torch.cuda.current_stream(self.device)
            self._side.wait_stream(cur)
            with torch.cuda.stream(self._side):
                full = self._exchange(blocks, B_local, pre)
            # pull mode: the local block starts when the rendezvous kernel has finished.  Launched before it, the
            # local block fills every SM and the rendezvous (one small CTA that the pull waits for — the critical
            # path at 8 GPUs) took 0.12-0.17 ms to get scheduled and to get its flag traffic through
            # (profiles/r02_aq_timeline_n8.jsonl)
            if self._after_barrier is not None and self._local_after_barrier(blocks, B_local):
                cur.wait_event(self._after_barrier)
            self._mul_local(blocks, B_local, out)          # overlaps the exchange
            cur.wait_stream(self._side)
            full.record_stream(cur)
        else:
            full = self._exchange(blocks, B_local, pre)
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
                 weight_decay=5e-4, seed=15, association="reference"):
        super().__init__()
        from .layers import GraphConvolution, GraphConvolution2
        self.part = part
        self.association = association   # see gcn_b200.model: 'auto' = SpMM at the narrower width
        self._ax = None
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
        logits, nc = self.forward_logits(x_local)
        return torch.nn.functional.log_softmax(logits[:, :nc] if logits.shape[1] != nc else logits, dim=1)

    def forward_logits(self, x_local):
        """(logits, nclass) of this rank's rows; `logits` may carry zero padding columns past nclass."""
        from . import gemm
        F = torch.nn.functional
        nfeat, nhid = self.gc1.weight.shape
        nclass = self.gc2.weight.shape[1]
        if (self.association == "auto" and nfeat <= nhid and nclass < nhid
                and not torch.cuda.is_current_stream_capturing()):
            # fused dense block between the cached Â·X and the layer-2 aggregation (gcn_b200.gemm.mlp2)
            if self._ax is None or self._ax[0] is not x_local or self._ax[1] != x_local._version:
                with torch.no_grad():
                    self._ax = (x_local, x_local._version, self.part.forward(x_local))
            pad = (-nclass) % 4
            w = F.pad(self.gc2.weight, (0, pad)) if pad else self.gc2.weight
            self._drop_step = getattr(self, "_drop_step", 0) + 1
            seed = (torch.initial_seed() * 2654435761 + 7919 * self.part.rank + self._drop_step) & 0xFFFFFFFF
            s2 = gemm.mlp2(self._ax[2], self.gc1.weight, self.gc1.bias, w, p=self.dropout, seed=seed,
                           training=self.training, out=self.part.next_operand(nclass + pad))
            b2 = F.pad(self.gc2.bias, (0, pad)) if pad else self.gc2.bias
            return spmm(self.part, s2) + b2, nclass
        if self.association == "auto" and nfeat <= nhid:
            if self._ax is None or self._ax[0] is not x_local or self._ax[1] != x_local._version:
                with torch.no_grad():   # Â·X once: X is constant over the epochs
                    self._ax = (x_local, x_local._version, self.part.forward(x_local))
            h = gemm.xw(self._ax[2], self.gc1.weight, bias=self.gc1.bias, relu=True)
        else:
            support = gemm.xw(x_local, self.gc1.weight, out=self.part.next_operand(nhid))   # straight into the exchange slot
            h = torch.relu(spmm(self.part, support) + self.gc1.bias)
        h = F.dropout(h, self.dropout, training=self.training)
        if self.association == "auto" and nclass < nhid:
            pad = (-nclass) % 4
            w = F.pad(self.gc2.weight, (0, pad)) if pad else self.gc2.weight
            s2 = gemm.xw(h, w, out=self.part.next_operand(nclass + pad))
            b2 = F.pad(self.gc2.bias, (0, pad)) if pad else self.gc2.bias
            out = spmm(self.part, s2) + b2
        else:
            s2 = spmm(self.part, h)
            out = gemm.xw(s2, self.gc2.weight, bias=self.gc2.bias)
        return out, nclass

    def step(self, x_local, labels_local, idx_train_local, n_train_global):
        """One training epoch (forward, NLL on the local training rows, backward, gradient
        all-reduce, Adam).  Returns the local loss contribution (sum / global count)."""
        self.train()
        self.optimizer.zero_grad(set_to_none=True)
        logits, nc = self.forward_logits(x_local)
        if logits.is_cuda:   # fused log-softmax + NLL (ops.logsoftmax_nll); the divisor is the GLOBAL training count
            cache = getattr(self, "_train_mask", None)
            if cache is None or cache[0] is not idx_train_local:
                m = torch.zeros(logits.shape[0], dtype=torch.uint8, device=logits.device)
                m[idx_train_local] = 1
                cache = self._train_mask = (idx_train_local, m)
            _, loss = ops.logsoftmax_nll(logits.contiguous(), nc, labels_local, idx_train_local, is_train=cache[1],
                                         denom=n_train_global)
        else:
            out = torch.nn.functional.log_softmax(logits[:, :nc], dim=1)
            loss = torch.nn.functional.nll_loss(out[idx_train_local], labels_local[idx_train_local],
                                                reduction="sum") / n_train_global
        loss.backward()
        allreduce_grads(self, self.part.group)
        self.optimizer.step()
        return loss

"""2-layer GCN with the reference's constructor / ``fit`` / ``forward`` / ``test`` / ``predict``
API (pygcn/gcn5.py:146-411, flex form gcn6.py:201-557), running on the B200 SpMM engine.

The immediate caller of the hot path.  Kept from the reference: argument
names and defaults, Adam(lr, weight_decay) over all parameters, NLL loss on ``idx_train``, the
three training loops (without validation / with validation / early stopping), and the per-dataset
layer-2 choice (A·(XW) for pubmed/flickr, (A·X)·W otherwise, gcn5.py:160-163).  Changed on purpose:

  * ``fit`` accepts an optional vertex reordering (``reorder='rabbit'|'gorder'|'rcm'|'deg'|'dfs'|None``,
    the module constants OVO/DFS/GOR/RBT of gcn6.py:27-30): Â is renumbered on the host
    (bit-exact with the reference's renumber.cu), features AND labels AND index sets are permuted
    consistently (the reference permutes idx_train but not the labels, SURVEY.md §4 defect 2).
  * backward uses Âᵀ; ``_train_with_val`` / ``predict`` work with the same forward signature
    (they are broken in gcn6, SURVEY.md §4 defect 4).
  * ReLU of layer 1 is fused into the producing kernel's epilogue.
  * ``association='auto'`` (constructor; default ``'reference'``) evaluates the same two layers
    ``Â·X·W + b`` in the cheaper of the two orders the reference itself offers as GraphConvolution
    (Â·(X·W)) and GraphConvolution2 ((Â·X)·W) — cf. gcn6.py:215-218 — so that the SpMM always runs at
    the narrower width: layer 1 with nfeat <= nhid becomes (Â·X)·W1 with Â·X computed ONCE per fit
    (X is constant, so neither the forward nor the backward of layer 1 contains an SpMM any more),
    layer 2 with nclass < nhid becomes Â·(H·W2) (SpMM at width nclass, padded to a multiple of 4,
    instead of nhid).  Same mathematics, different rounding order: outputs agree with the
    reference association to ~1e-6 relative (tests/test_layers_gpu.py).
"""
from __future__ import annotations

from copy import deepcopy

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
import torch.optim as optim

from . import gemm, graph, ops
from .layers import GraphConvolution, GraphConvolution2, _is_sparse


def accuracy(output, labels):
    """utils.accuracy (pygcn/gcnio/util/utils.py)."""
    preds = output.max(1)[1].type_as(labels)
    return preds.eq(labels).double().sum() / len(labels)


class GCN(nn.Module):
    def __init__(self, nfeat, nhid, nclass, dataset="dataset", dropout=0.5, lr=0.01,
                 weight_decay=5e-4, with_relu=True, with_bias=True, device=None,
                 association="reference"):
        super().__init__()
        assert device is not None, "Please specify 'device'!"
        if association not in ("reference", "auto"):
            raise ValueError("association must be 'reference' or 'auto'")
        self.association = association
        self._ax_cache = None   # (features tensor, its _version, adjacency object, Â·X): strong references
        self.device = torch.device(device)
        self.nfeat = nfeat
        self.hidden_sizes = [nhid]
        self.nclass = nclass
        self.dataname = dataset
        self.gc1 = GraphConvolution(nfeat, nhid, with_bias=with_bias, name=dataset, layer="layer1")
        if dataset in ("pubmed", "flickr"):
            self.gc2 = GraphConvolution(nhid, nclass, with_bias=with_bias, name=dataset, layer="layer2")
        else:
            self.gc2 = GraphConvolution2(nhid, nclass, with_bias=with_bias, name=dataset, layer="layer2")
        self.dropout = dropout
        self.lr = lr
        self.weight_decay = weight_decay if with_relu else 0
        self.with_relu = with_relu
        self.with_bias = with_bias
        self.output = None
        self.adj_norm = None
        self.features = None
        self.labels = None
        self.vomp = None  # vomp[new] = old when a reordering is active
        self.to(self.device)

    def reset_timing(self):
        pass

    def forward(self, x, adj, name="dataset"):
        """adj: normalised adjacency (torch sparse tensor on the GPU or an SpmmPlan)."""
        logits, nc = self.forward_logits(x, adj)
        return F.log_softmax(logits[:, :nc] if logits.shape[1] != nc else logits, dim=1)

    def forward_logits(self, x, adj):
        """(logits, nclass): the class scores before log-softmax; `logits` may carry zero padding columns
        past nclass (the aggregation runs at a multiple of 4 columns)."""
        if self.association == "auto":
            return self._forward_min_width(x, adj)
        x = self.gc1(x, adj, relu=self.with_relu)
        x = F.dropout(x, self.dropout, training=self.training)
        x = self.gc2(x, adj)
        return x, x.shape[1]

    def invalidate_cache(self):
        """Forget the cached Â·X (call after changing the features in place)."""
        self._ax_cache = None

    def _aggregated_features(self, x, adj):
        # the cache holds the tensor and the adjacency themselves (an address or id() can be re-used by a
        # new object once the old one is freed) and the tensor's version counter (in-place edits)
        c = self._ax_cache
        if c is None or c[0] is not x or c[1] != x._version or c[2] is not adj:
            with torch.no_grad():
                self._ax_cache = (x, x._version, adj, ops.plan_for(adj).spmm(x))
        return self._ax_cache[3]

    def _forward_min_width(self, x, adj):
        nfeat, nhid = self.gc1.weight.shape
        nh, nc = self.gc2.weight.shape
        if (not _is_sparse(x) and not x.requires_grad and nfeat <= nhid and nc < nh and self.with_relu
                and not torch.cuda.is_current_stream_capturing()):
            # both layers in the min-width association: everything between the cached Â·X and the layer-2
            # aggregation is ONE fused block — (Â·X)·W1 + b1, ReLU, dropout in the first product's epilogue,
            # ·W2 next, dropout/ReLU backward in the epilogue of dS·W2ᵀ (gcn_b200.gemm.mlp2).  Under CUDA-graph
            # capture the seed would be frozen into the graph, so capture takes the unfused path below.
            pad = (-nc) % 4
            w2 = F.pad(self.gc2.weight, (0, pad)) if pad else self.gc2.weight
            b2 = self.gc2.bias
            if b2 is not None and pad:
                b2 = F.pad(b2, (0, pad))
            self._drop_step = getattr(self, "_drop_step", 0) + 1
            seed = (torch.initial_seed() * 2654435761 + self._drop_step) & 0xFFFFFFFF
            s2 = gemm.mlp2(self._aggregated_features(x, adj), self.gc1.weight, self.gc1.bias, w2,
                           p=self.dropout, seed=seed, training=self.training)
            return ops.spmm(adj, s2, bias=b2), nc
        if not _is_sparse(x) and not x.requires_grad and nfeat <= nhid:
            # (Â·X)·W1 + b1 with Â·X cached: no SpMM in layer 1, forward or backward
            h = gemm.xw(self._aggregated_features(x, adj), self.gc1.weight, bias=self.gc1.bias,
                        relu=self.with_relu)
        else:
            h = self.gc1(x, adj, relu=self.with_relu) if isinstance(self.gc1, GraphConvolution) else self.gc1(x, adj)
        h = F.dropout(h, self.dropout, training=self.training)
        if nc < nh:
            # Â·(H·W2) + b2: SpMM at width nclass (zero-padded to 4 floats for 128-bit rows)
            pad = (-nc) % 4
            w = F.pad(self.gc2.weight, (0, pad)) if pad else self.gc2.weight
            b = self.gc2.bias
            if b is not None and pad:
                b = F.pad(b, (0, pad))
            out = ops.spmm(adj, gemm.xw(h, w), bias=b)
        else:
            out = self.gc2(h, adj)
        return out, nc

    def initialize(self):
        self.gc1.reset_parameters()
        self.gc2.reset_parameters()

    # ---- data preparation ----------------------------------------------------------------------
    def _prepare(self, features, adj, labels, normalize, reorder):
        """-> (features on device (dense tensor or SpmmPlan), adjacency SpmmPlan, labels, vomp)."""
        import scipy.sparse as sp

        if isinstance(adj, ops.SpmmPlan):
            if reorder:
                raise ValueError("reorder needs the adjacency itself, not a prebuilt plan")
            plan, vomp = adj, None
        else:
            if isinstance(adj, torch.Tensor):
                a = adj.coalesce() if adj.layout == torch.sparse_coo else adj.to_sparse_coo().coalesce()
                idx = a.indices().cpu().numpy()
                mx = sp.csr_matrix((a.values().cpu().numpy(), (idx[0], idx[1])), shape=tuple(a.shape))
            else:
                mx = sp.csr_matrix(adj)
            mx = graph.normalize_adj(mx) if normalize else mx.astype(np.float32)
            mx.sort_indices()
            vomp = None
            rp, ci, va = mx.indptr, mx.indices, mx.data
            if reorder:
                fn = {"gorder": graph.renumber_gorder, "rcm": graph.renumber_rcm, "rabbit": graph.renumber_rabbit,
                      "deg": graph.renumber_deg, "dfs": graph.renumber_dfs}[reorder]
                rp, ci, va, vomp = fn(rp, ci, va)
            plan = ops.SpmmPlan(rp, ci, va, mx.shape, device=self.device)

        if isinstance(features, torch.Tensor) and not features.is_sparse:
            x = features.to(self.device, torch.float32)
            if vomp is not None:
                x = ops.permute_rows(x, torch.from_numpy(vomp).to(self.device))
        elif isinstance(features, ops.SpmmPlan):
            if vomp is not None:
                raise ValueError("reorder with prebuilt feature plan")
            x = features
        else:
            if isinstance(features, torch.Tensor):
                f = features.coalesce()
                idx = f.indices().cpu().numpy()
                fx = sp.csr_matrix((f.values().cpu().numpy(), (idx[0], idx[1])), shape=tuple(f.shape))
            else:
                fx = sp.csr_matrix(features)
            if vomp is not None:
                fx = fx[vomp]
            x = ops.SpmmPlan.from_scipy(fx.astype(np.float32), device=self.device)

        y = torch.as_tensor(np.asarray(labels) if not isinstance(labels, torch.Tensor) else labels)
        y = y.to(self.device, torch.long)
        if vomp is not None:
            y = y[torch.from_numpy(vomp.astype(np.int64)).to(self.device)]
        return x, plan, y, vomp

    def _map_index(self, idx):
        """Old vertex ids -> positions after reordering (permutateIdx, gcn6.py:378)."""
        if idx is None:
            return None
        idx = torch.as_tensor(np.asarray(idx) if not isinstance(idx, torch.Tensor) else idx).long()
        if self.vomp is not None:
            rank = np.empty(len(self.vomp), dtype=np.int64)
            rank[self.vomp] = np.arange(len(self.vomp))
            idx = torch.from_numpy(rank)[idx.cpu()]
        return idx.to(self.device)

    def fit(self, features, adj, labels, idx_train, idx_val=None, train_iters=200, initialize=True,
            verbose=False, normalize=True, patience=500, name="dataset", reorder=None, cuda_graph=False):
        self.device = self.gc1.weight.device
        if self.device.type != "cuda":
            raise RuntimeError("gcn_b200: GCN must live on a CUDA device (no CPU fallback)")
        if initialize:
            self.initialize()
        self._ax_cache = None
        self.features, self.adj_norm, self.labels, self.vomp = self._prepare(
            features, adj, labels, normalize, reorder)
        idx_train = self._map_index(idx_train)
        idx_val = self._map_index(idx_val)
        if idx_val is None:
            self._train_without_val(self.labels, idx_train, train_iters, verbose, cuda_graph=cuda_graph)
        elif patience < train_iters:
            self._train_with_early_stopping(self.labels, idx_train, idx_val, train_iters, patience, verbose)
        else:
            self._train_with_val(self.labels, idx_train, idx_val, train_iters, verbose)

    def _loss(self, labels, idx_train):
        """(output, loss) of one forward pass: log-softmax over all vertices + NLL on idx_train, fused
        (ops.logsoftmax_nll: gcn6.py:248 + :427 in one pass forward, one pass backward)."""
        logits, nc = self.forward_logits(self.features, self.adj_norm)
        if logits.stride(1) != 1:
            logits = logits.contiguous()
        cache = getattr(self, "_train_mask", None)
        if cache is None or cache[0] is not idx_train or cache[1].shape[0] != logits.shape[0]:
            m = torch.zeros(logits.shape[0], dtype=torch.uint8, device=logits.device)
            m[idx_train] = 1
            cache = self._train_mask = (idx_train, m)
        return ops.logsoftmax_nll(logits, nc, labels, idx_train, is_train=cache[1])

    def _step(self, optimizer, labels, idx_train):
        self.train()
        optimizer.zero_grad()
        output, loss = self._loss(labels, idx_train)
        loss.backward()
        optimizer.step()
        return output, loss

    def graphed_step(self, optimizer, labels, idx_train, warmup=3):
        """One training step (forward, NLL, backward, optimizer) captured in a CUDA graph.

        The epoch of a small graph is launch-bound (Cora-shaped: ~60 kernels, 1.4 ms eager); the
        kernels of libgcn_b200.so take torch's current stream and use only stream-ordered
        allocations, so they capture like torch's own.  `optimizer` must be capturable
        (``Adam(..., capturable=True)``).  `warmup` eager steps run first (they are real training
        steps).  Returns ``replay() -> (output, loss)``: static tensors refreshed by every replay."""
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._step(optimizer, labels, idx_train)
        torch.cuda.current_stream(self.device).wait_stream(side)
        optimizer.zero_grad(set_to_none=True)
        graph_ = torch.cuda.CUDAGraph()
        self.train()
        with torch.cuda.graph(graph_):
            output, loss = self._loss(labels, idx_train)
            loss.backward()
            optimizer.step()

        def replay():
            graph_.replay()
            return output, loss
        replay.graph = graph_
        return replay

    def _train_without_val(self, labels, idx_train, train_iters, verbose, cuda_graph=False):
        if cuda_graph and train_iters > 3:
            optimizer = optim.Adam(self.parameters(), lr=self.lr, weight_decay=self.weight_decay, capturable=True)
            replay = self.graphed_step(optimizer, labels, idx_train, warmup=3)
            output = None
            for i in range(3, train_iters):
                output, loss = replay()
                if verbose and i % 10 == 0:
                    print(f"Epoch {i:3d}, training loss: {loss.item():.6f}")
            self.output = output
            return
        optimizer = optim.Adam(self.parameters(), lr=self.lr, weight_decay=self.weight_decay)
        output = None
        for i in range(train_iters):
            output, loss = self._step(optimizer, labels, idx_train)
            if verbose and i % 10 == 0:
                print(f"Epoch {i:3d}, training loss: {loss.item():.6f}")
        self.output = output

    def _train_with_val(self, labels, idx_train, idx_val, train_iters, verbose):
        optimizer = optim.Adam(self.parameters(), lr=self.lr, weight_decay=self.weight_decay)
        best_loss_val, best_acc_val = 100, 0
        weights = deepcopy(self.state_dict())
        for i in range(train_iters):
            _, loss = self._step(optimizer, labels, idx_train)
            if verbose and i % 10 == 0:
                print(f"Epoch {i}, training loss: {loss.item()}")
            self.eval()
            with torch.no_grad():
                output = self.forward(self.features, self.adj_norm)
            loss_val = F.nll_loss(output[idx_val], labels[idx_val])
            acc_val = accuracy(output[idx_val], labels[idx_val])
            if best_loss_val > loss_val:
                best_loss_val = loss_val
                self.output = output
                weights = deepcopy(self.state_dict())
            if acc_val > best_acc_val:
                best_acc_val = acc_val
                self.output = output
                weights = deepcopy(self.state_dict())
        self.load_state_dict(weights)

    def _train_with_early_stopping(self, labels, idx_train, idx_val, train_iters, patience, verbose):
        optimizer = optim.Adam(self.parameters(), lr=self.lr, weight_decay=self.weight_decay)
        early_stopping = patience
        best_loss_val = 100
        weights = deepcopy(self.state_dict())
        for i in range(train_iters):
            _, loss = self._step(optimizer, labels, idx_train)
            if verbose and i % 10 == 0:
                print(f"Epoch {i}, training loss: {loss.item()}")
            self.eval()
            with torch.no_grad():
                output = self.forward(self.features, self.adj_norm)
            loss_val = F.nll_loss(output[idx_val], labels[idx_val])
            if best_loss_val > loss_val:
                best_loss_val = loss_val
                self.output = output
                weights = deepcopy(self.state_dict())
                patience = early_stopping
            else:
                patience -= 1
            if i > early_stopping and patience <= 0:
                break
        self.load_state_dict(weights)

    def test(self, idx_test):
        self.eval()
        with torch.no_grad():
            output = self.predict()
        idx_test = self._map_index(idx_test)
        loss_test = F.nll_loss(output[idx_test], self.labels[idx_test])
        acc_test = accuracy(output[idx_test], self.labels[idx_test])
        print(f"Test set results: loss= {loss_test.item():.4f} accuracy= {acc_test.item():.4f}")
        return acc_test

    def predict(self, features=None, adj=None):
        """Log-probabilities in the (possibly reordered) vertex numbering of the last ``fit``."""
        self.eval()
        if features is not None or adj is not None:
            self._ax_cache = None
            labels = self.labels if self.labels is not None else np.zeros(features.shape[0], dtype=np.int64)
            self.features, self.adj_norm, _, self.vomp = self._prepare(features, adj, labels, True, None)
        return self.forward(self.features, self.adj_norm)

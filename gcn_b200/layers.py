"""GraphConvolution layers with the reference's torch API, computed by the B200 SpMM engine.

Same class names, constructor arguments, parameter names/shapes/initialisation and
``forward(input, adj)`` signature as pygcn/gcn5.py:21-139 (the torch-API form of the layer; gcn1.py:
22-58 is the original).  What changes is what runs underneath:

  * ``torch.spmm(adj, support)`` (gcn5.py:90, :135)  ->  :func:`gcn_b200.ops.spmm` (hand-written
    sm_100a kernels behind the C ABI), with the ``+ bias`` of gcn5.py:92 fused into the kernel
    epilogue for the A·(XW) layer;
  * ``torch.spmm(input, W)`` for sparse features (gcn5.py:85) -> the same SpMM engine with the
    feature matrix as the sparse operand (its plan is cached on the tensor);
  * ``torch.mm(input, W)`` (gcn5.py:88) -> :func:`gcn_b200.gemm.xw` (tcgen05 tensor cores).

``adj`` is a torch sparse COO/CSR tensor on the GPU (what the reference passes) or a prebuilt
:class:`gcn_b200.ops.SpmmPlan`.  The per-layer timers (dmk) and the constructor's print are the
reference's instrumentation, not part of the operator, and are not reproduced.
"""
from __future__ import annotations

import math

import torch
from torch.nn.modules.module import Module
from torch.nn.parameter import Parameter

from . import gemm, ops


def _is_sparse(t) -> bool:
    return isinstance(t, ops.SpmmPlan) or (isinstance(t, torch.Tensor)
                                           and (t.is_sparse or t.layout == torch.sparse_csr))


class _GraphConvBase(Module):
    def __init__(self, in_features, out_features, with_bias=True, name="dataset", layer="layer0"):
        super().__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.weight = Parameter(torch.empty(in_features, out_features, dtype=torch.float32))
        self.layer = layer
        self.name = name
        if with_bias:
            self.bias = Parameter(torch.empty(out_features, dtype=torch.float32))
        else:
            self.register_parameter("bias", None)
        self.reset_parameters()

    def reset_parameters(self):
        # uniform(±1/sqrt(out_features)) for weight and bias — pygcn/gcn5.py:41-50
        stdv = 1.0 / math.sqrt(self.weight.size(1))
        self.weight.data.uniform_(-stdv, stdv)
        if self.bias is not None:
            self.bias.data.uniform_(-stdv, stdv)

    def reset_timing(self):  # kept for call-compatibility with GCN.reset_timing (gcn5.py:179-181)
        pass

    def __repr__(self):
        return f"{self.__class__.__name__} ({self.in_features} -> {self.out_features})"


class GraphConvolution(_GraphConvBase):
    """A·(X·W) + bias   — pygcn/gcn5.py:21-99 (flex form: gcn6.py:66-148)."""

    def forward(self, input, adj, relu=False):
        if _is_sparse(input):
            # sparse features: X·W is itself an SpMM (gcn5.py:85); gradient flows to W only
            support = ops.spmm(input, self.weight)
        else:
            support = gemm.xw(input, self.weight)
        # bias (gcn5.py:91-92) and, on request, the ReLU that GCN.forward applies next (gcn5.py:191)
        # ride in the SpMM epilogue
        return ops.spmm(adj, support, bias=self.bias, relu=relu)


class GraphConvolution2(_GraphConvBase):
    """(A·X)·W + bias   — pygcn/gcn5.py:101-146 (flex form: gcn6.py:151-199)."""

    def forward(self, input, adj, relu=False):
        support = ops.spmm(adj, input)
        return gemm.xw(support, self.weight, bias=self.bias, relu=relu)

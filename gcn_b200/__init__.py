"""gcn_b200 — B200-native GCN aggregation path (SpMM Â·B / Âᵀ·dY behind pygcn's GraphConvolution).

Host side of the C ABI in ``include/gcn_b200.h`` (``libgcn_b200.so``, hand-written sm_100a CUDA).
Modules: ``ops`` (SpMM operator, plans), ``gemm`` (tcgen05 X·W), ``layers`` / ``model`` (the reference's
GraphConvolution / GCN API), ``dist`` (row partition over the GPUs of one box, NVLink halo pull),
``graph`` (bit-exact reorderings, renumbering, reference tile format), ``data`` (dataset files),
``synth`` (synthetic graphs of the BASELINE shapes).  CUDA only: importing ``_ffi`` fails if the
library is not built, and nothing falls back to the CPU.
"""
__version__ = "0.1.0"

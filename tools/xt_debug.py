"""Which MN-major operand layout does the tensor core accept for tf32?  Runs dW = X^T·G on a small case for the
GCNB_XT_VARIANT given in the environment and prints the error against fp64 and a few entries."""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gcn_b200 import gemm
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(3)
for (K, M, N, prec) in [(8192, 100, 64, 1), (8192, 100, 64, 0), (8192, 128, 256, 0), (50000, 48, 256, 0), (8192, 256, 48, 0)]:
    X = torch.randn(K, M, device=dev, generator=g)
    G = torch.randn(K, N, device=dev, generator=g)
    ref = X.double().t() @ G.double()
    out = gemm.gemm_raw(X, G, trans_a=True, precision=prec)
    err = float((out.double() - ref).abs().max() / ref.abs().max())
    print(json.dumps({"variant": os.environ.get("GCNB_XT_VARIANT", "0"), "K": K, "M": M, "N": N, "precision": prec, "rel_err": err,
                      "out00": float(out[0, 0]), "ref00": float(ref[0, 0]), "out_absmax": float(out.abs().max()),
                      "ref_absmax": float(ref.abs().max())}), flush=True)

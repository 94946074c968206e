"""Times dW = Xᵀ·dS (gcnb_xw, tall K) at the products-shaped epoch's sizes: the prepared-image path
(GCNB_XW_TALLK=0) against the in-kernel conversion of the second operand (=1) over the split-chain
bound (GCNB_XW_MAX_CHAIN, k-blocks of 16) and the n-tile width; error against an fp64 product.
One JSON line per point."""
import json
import os
import sys

import torch

from gcn_b200 import gemm

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_449_029
dev = "cuda:0"
wide = torch.randn(n, 256, device=dev)
mid = torch.randn(n, 100, device=dev)
thin = torch.randn(n, 48, device=dev)[:, :47]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(A, B, env):
    for k in ("GCNB_XW_TALLK", "GCNB_XW_MAX_CHAIN", "GCNB_XW_TALLK_BN"):
        os.environ.pop(k, None)
    os.environ.update(env)
    ts = []
    for it in range(5):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = gemm.gemm_raw(A, B, trans_a=True)
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2], out


points = [("dW1 swapped: dZ1[n,256]^T . AX[n,100]", wide, mid), ("dW2: H[n,256]^T . dS2[n,47]", wide, thin)]
variants = [("image", {"GCNB_XW_TALLK": "0"}), ("tallk default", {}),
            ("chain32", {"GCNB_XW_MAX_CHAIN": "32"}), ("chain128", {"GCNB_XW_MAX_CHAIN": "128"}),
            ("chain512", {"GCNB_XW_MAX_CHAIN": "512"}), ("bn64", {"GCNB_XW_TALLK_BN": "64"}),
            ("bn64 chain128", {"GCNB_XW_TALLK_BN": "64", "GCNB_XW_MAX_CHAIN": "128"})]
for name, A, B in points:
    ref = A.double().t() @ B.double()
    gb = (A.numel() + B.numel() + ref.numel()) * 4 / 1e9
    for vname, env in variants:
        ms, out = timed(A, B, env)
        err = float((out.double() - ref).abs().max() / ref.abs().max())
        print(json.dumps({"point": name, "variant": vname, "n": n, "ms": round(ms, 3), "algorithmic_GB": round(gb, 3),
                          "GBps": round(gb / ms * 1e3, 1), "rel_err_fp64": err}), flush=True)
    ms, out = timed(A, B, {})
    t = torch.mm(A.t(), B)
    print(json.dumps({"point": name, "variant": "torch.mm fp32 (error only)",
                      "rel_err_fp64": float((t.double() - ref).abs().max() / ref.abs().max())}), flush=True)

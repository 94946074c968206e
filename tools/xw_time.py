"""Device time of the dense products of the products-shaped epoch, one by one (CUDA events, median of 7).
    python tools/xw_time.py [n]
Prints one JSON line per product: shape, kernel route, ms, GB moved (operands once), GB/s."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gcn_b200 import gemm  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_449_029
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
AX = torch.randn(n, 100, device=dev, generator=g)
W1 = torch.randn(100, 256, device=dev, generator=g) / 10
b1 = torch.randn(256, device=dev, generator=g)
W2 = torch.randn(256, 48, device=dev, generator=g) / 16
bits = torch.empty(n, 8, dtype=torch.int32, device=dev)
h = torch.empty(n, 256, device=dev)
s2 = torch.empty(n, 48, device=dev)
gs = torch.randn(n, 48, device=dev, generator=g)
dpre = torch.empty(n, 256, device=dev)


def timed(fn, reps=7):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


cases = [
    ("fwd1 h = dropout(relu(AX·W1+b1))  M=n K=100 N=256", lambda: gemm.gemm_raw(AX, W1, bias=b1, relu=True, dropout=(0.5, 7), mask_out=bits, out=h), 4 * n * (100 + 256)),
    ("fwd2 s2 = h·W2                    M=n K=256 N=48", lambda: gemm.gemm_raw(h, W2, out=s2), 4 * n * (256 + 48)),
    ("dpre = (g·W2^T) masked            M=n K=48  N=256", lambda: gemm.gemm_raw(gs, W2, trans_b=True, mask=(bits, 0.5), out=dpre), 4 * n * (48 + 256)),
    ("dW2 = h^T·g                       K=n M=256 N=48", lambda: gemm._dw(h, gs, 0), 4 * n * (256 + 48)),
    ("dW1 = AX^T·dpre                   K=n M=100 N=256", lambda: gemm._dw(AX, dpre, 0), 4 * n * (100 + 256)),
]
if os.environ.get("XW_TIME_VARIANTS"):
    # which part of the layer-1 product costs what: epilogue pieces and the 3xTF32 machinery, one at a time
    cases = [
        ("fwd1 plain        AX·W1", lambda: gemm.gemm_raw(AX, W1, out=h), 4 * n * 356),
        ("fwd1 +bias", lambda: gemm.gemm_raw(AX, W1, bias=b1, out=h), 4 * n * 356),
        ("fwd1 +relu", lambda: gemm.gemm_raw(AX, W1, relu=True, out=h), 4 * n * 356),
        ("fwd1 +dropout", lambda: gemm.gemm_raw(AX, W1, dropout=(0.5, 7), out=h), 4 * n * 356),
        ("fwd1 +mask_out", lambda: gemm.gemm_raw(AX, W1, mask_out=bits, out=h), 4 * n * 356),
        ("fwd1 all", lambda: gemm.gemm_raw(AX, W1, bias=b1, relu=True, dropout=(0.5, 7), mask_out=bits, out=h), 4 * n * 356),
        ("fwd1 plain, single-pass TF32", lambda: gemm.gemm_raw(AX, W1, out=h, precision=1), 4 * n * 356),
        ("dpre plain        g·W2^T", lambda: gemm.gemm_raw(gs, W2, trans_b=True, out=dpre), 4 * n * 304),
        ("dpre masked", lambda: gemm.gemm_raw(gs, W2, trans_b=True, mask=(bits, 0.5), out=dpre), 4 * n * 304),
        ("dpre plain, single-pass TF32", lambda: gemm.gemm_raw(gs, W2, trans_b=True, out=dpre, precision=1), 4 * n * 304),
        ("copy h -> dpre (2.5 GB in, 2.5 GB out)", lambda: dpre.copy_(h), 8 * n * 256),
        ("fill dpre (2.5 GB out)", lambda: dpre.fill_(1.0), 4 * n * 256),
    ]
only = os.environ.get("XW_TIME_ONLY")   # substring filter on the product names
for name, fn, nbytes in cases:
    if only and only not in name:
        continue
    ms = timed(fn)
    print(json.dumps({"product": name, "ms": round(ms, 4), "GB": round(nbytes / 1e9, 3), "GBps": round(nbytes / ms / 1e6, 1),
                      "env": {k: v for k, v in os.environ.items() if k.startswith("GCNB_X")}}), flush=True)

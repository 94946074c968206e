"""Reorder sweep on the Reddit-shaped graph at its stated size (233 k vertices, 114.8 M nnz — BASELINE config 3):
identity / degree / RCM / Gorder(3) / Rabbit with the permutations cached in profiles/perm/ (host seconds in
profiles/perm/reddit_host_seconds_r02.json), k = 128 and 602, stream class and (k = 128) the opt-in panel class.
One graph generation (CPU generator, the one the cached permutations belong to), one JSON line per point.
The first version of this tool recorded the start event on an idle GPU, so host time of the call landed inside
the interval (profiles/r02_n_reddit_sweep.jsonl: up to +2 ms); it now times like bench.py.
    python tools/reddit_sweep.py [orders...]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gcn_b200 import graph, ops, synth  # noqa: E402

dev = torch.device("cuda", 0)
orders = sys.argv[1:] or ["identity", "deg", "rcm", "gorder", "rabbit"]
t0 = time.perf_counter()
rp, ci, va, n = synth.make_graph("reddit", device="cpu")
nnz = int(rp[-1])
print(json.dumps({"graph": "reddit-shaped", "n": n, "nnz": nnz, "generate_s": round(time.perf_counter() - t0, 1)}), flush=True)
host = json.load(open(os.path.join(ROOT, "profiles", "perm", "reddit_host_seconds_r02.json")))


import bench  # noqa: E402  (timing as in bench.py: launches queued behind a spin kernel, events on the launching stream)


def timed(fn, reps=10):
    return bench._event_ms(fn, reps=reps, warm=3)


g = torch.Generator(device=dev).manual_seed(15)
B = {k: torch.randn(n, k, device=dev, generator=g) for k in (128, 602)}
for o in orders:
    vomp = np.load(os.path.join(ROOT, "profiles", "perm", f"reddit_{o}_vomp.npy")).astype(np.int32)
    r2, c2, v2 = (rp, ci, va) if o == "identity" else graph.perm_apply(rp, ci, va, vomp)
    for panel in (False, True):
        plan = ops.SpmmPlan(r2, c2, v2, (n, n), device=dev, assume_symmetric=True, panel=panel)
        info = plan.info()
        for k in ((128, 602) if not panel else (128,)):
            C = torch.empty(n, k, device=dev)
            ms = timed(lambda: plan.spmm(B[k], out=C))
            abytes, flops = synth.spmm_bytes_flops(n, n, nnz, k)
            # checksum of checksums against the host CSR (the permuted operand is a plain random matrix here)
            colsum = torch.from_numpy(np.bincount(c2, weights=v2.astype(np.float64), minlength=n)).to(dev)
            chk = float(((C.double().sum(0) - colsum @ B[k].double()).abs().max() / (colsum @ B[k].double()).abs().max()))
            print(json.dumps({"order": o, "host_order_s": host.get(o, 0), "class": "panel" if panel else "stream", "k": k,
                              "ms_median": round(ms, 4), "gflops": round(flops / ms / 1e6, 1),
                              "hbm_frac_algorithmic": round(abytes / ms / 1e6 / 6546.2, 4),
                              "nnz_per_distinct_panel_column": (round(nnz / info["panel_distinct_columns"], 3)
                                                                if info["panel_distinct_columns"] else None),
                              "checksum_rel_err": chk}), flush=True)
            del C
        del plan
        torch.cuda.empty_cache()

"""One rank's blocks of the row-partitioned SpMM, timed on ONE GPU (no peers): how long do the local-column block
and the accumulating remote-column block of rank r of P take by themselves, and what does the remote block look like
(rows touched, non-zeros per row, halo rows)?  ncu can be pointed at this (single process).
  python tools/remote_block_probe.py --parts 8 --rank 3 --k 100 [--seg-nnz N]"""
import argparse
import json
import os
import sys

ap = argparse.ArgumentParser()
ap.add_argument("--parts", type=int, default=8)
ap.add_argument("--rank", type=int, default=3)
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--seg-nnz", type=int, default=0)
ap.add_argument("--iters", type=int, default=20)
args = ap.parse_args()
if args.seg_nnz:
    os.environ["GCNB_SEG_NNZ"] = str(args.seg_nnz)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from gcn_b200 import dist as D, ops, synth

dev = torch.device("cuda:0")
rowptr, col, val, n = synth.make_graph("products", device=dev)
bounds = D.nnz_balanced_bounds(rowptr, args.parts)
local, remote, stride = D.split_local_remote(rowptr, col, val, bounds, args.rank)
remote_c, halo_part, halo_row = D.compact_halo(remote, bounds, stride)
m = int(bounds[args.rank + 1] - bounds[args.rank])
H = max(len(halo_row), 1)
rp = remote_c[0].astype(np.int64)
per_row = np.diff(rp)
info = {"parts": args.parts, "rank": args.rank, "k": args.k, "rows": m, "local_nnz": int(local[0][-1]),
        "remote_nnz": int(rp[-1]), "halo_rows": int(H), "rows_with_remote": int((per_row > 0).sum()),
        "remote_nnz_per_touched_row": float(rp[-1] / max(1, (per_row > 0).sum())),
        "max_remote_row": int(per_row.max()), "seg_nnz": args.seg_nnz}
p_local = ops.SpmmPlan(*local, (m, m), device=dev, need_transpose=False)
p_remote = ops.SpmmPlan(*remote_c, (m, H), device=dev, need_transpose=False)
g = torch.Generator(device=dev).manual_seed(1)
B = torch.randn(m, args.k, device=dev, generator=g)
Hb = torch.randn(H, args.k, device=dev, generator=g)
C = torch.empty(m, args.k, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn):
    ts = []
    for i in range(args.iters + 3):
        flush.zero_()                       # halo rows and C rows start in HBM, as after a pull / a local pass
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        if i >= 3:
            ts.append(a.elapsed_time(b))
    return float(np.median(ts)), float(min(ts))


info["local_ms"], info["local_ms_min"] = timed(lambda: p_local.spmm(B, out=C))
info["remote_accum_ms"], info["remote_accum_ms_min"] = timed(lambda: p_remote.spmm(Hb, out=C, accumulate=True))
info["remote_store_ms"], _ = timed(lambda: p_remote.spmm(Hb, out=C))
# check
ref = torch.zeros(m, args.k, device=dev, dtype=torch.float64)
C.zero_()
p_remote.spmm(Hb, out=C, accumulate=True)
rows = torch.from_numpy(np.repeat(np.arange(m), per_row)).to(dev)
cc = torch.from_numpy(remote_c[1].astype(np.int64)).to(dev)
vv = torch.from_numpy(remote_c[2]).to(dev).double()
ref.index_add_(0, rows, Hb.double()[cc] * vv[:, None])
info["rel_err"] = float((C.double() - ref).abs().max() / ref.abs().max())
print(json.dumps(info))

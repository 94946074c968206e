#!/usr/bin/env python
"""Top SASS instructions by stall samples from `ncu -i X.ncu-rep --page source --csv` (stdin or file)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
# several kernels may be concatenated: split on "Kernel Name" rows
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "data": []}
        blocks.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None and len(r) == len(cur["hdr"]):
        cur["data"].append(r)
for b in blocks[:1]:
    hdr, data = b["hdr"], b["data"]
    ia, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    tot = sum(int(r[isamp]) for r in data)
    print(b["name"][:100]); print("samples", tot, "warp-instructions", sum(int(r[iex]) for r in data), "SASS lines", len(data))
    sc = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    agg = {hdr[i]: sum(int(r[i] or 0) for r in data) for i in sc}
    print("stall totals:", ", ".join(f"{k[6:]} {100*v/tot:.1f}%" for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
    top = sorted(enumerate(data), key=lambda x: -int(x[1][isamp]))[:N]
    for idx, r in sorted(top):
        st = sorted(((hdr[i][6:], int(r[i] or 0)) for i in sc), key=lambda x: -x[1])[:2]
        print(f"{idx:5d} {r[ia].strip()[:64]:64s} {100*int(r[isamp])/tot:5.1f}% exec {r[iex]:>10s} {st}")

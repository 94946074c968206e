"""Timeline report of a GCNB_XS_TRACE file (clock64 stamps of CTA 0 of the streaming X·W kernel).
    python tools/xs_trace_report.py <trace file>
Roles: 0 producer issued chunk i, 1 converter saw chunk i (FULLA), 2 converter got a low-part slot (LOEMPTY),
3 converter arrived on CONV, 4 MMA thread passed its waits for chunk i, 5 MMA thread issued commits of chunk i,
6 epilogue warp 8 saw tile t complete (TFULL), 7 epilogue warp 8 finished tile t."""
import sys
import numpy as np

blocks, cur = [], None
for line in open(sys.argv[1]):
    if line.startswith("#"):
        cur = {"hdr": line[1:].strip(), "rows": {}}
        blocks.append(cur)
    elif cur is not None:
        v = line.split()
        cur["rows"][int(v[0])] = np.array([int(x) for x in v[1:]], dtype=np.int64)
for hdr in sorted({b["hdr"] for b in blocks}):
    b = [x for x in blocks if x["hdr"] == hdr][-1]
    r = b["rows"]
    t0 = r[0][0]
    print("==", hdr)
    nkc = int(hdr.split("nkc=")[1].split()[0])
    steady = slice(64, 448)
    for role, name in [(0, "producer issue"), (1, "conv sees X"), (3, "conv arrive"), (4, "MMA waits done"), (5, "MMA commits issued")]:
        d = np.diff(r[role][steady])
        print(f"  {name:20s} period/chunk: median {np.median(d):7.0f}  mean {d.mean():7.0f}  p90 {np.percentile(d, 90):7.0f} clk")
    nt = min(112, 448 // nkc)
    for role, name in [(6, "epilogue sees tile"), (7, "epilogue done")]:
        d = np.diff(r[role][16:nt])
        print(f"  {name:20s} period/tile : median {np.median(d):7.0f}  mean {d.mean():7.0f} clk  ({nkc} chunks per tile)")
    i = np.arange(64, 448)
    print(f"  X latency   (issue -> converter sees it)      median {np.median(r[1][i] - r[0][i]):7.0f} clk")
    print(f"  slot wait   (sees X -> has low-part slot)      median {np.median(r[2][i] - r[1][i]):7.0f} clk")
    print(f"  convert     (slot -> arrive)                   median {np.median(r[3][i] - r[2][i]):7.0f} clk")
    print(f"  conv -> MMA (arrive -> MMA thread past waits)  median {np.median(r[4][i] - r[3][i]):7.0f} clk")
    print(f"  MMA issue   (past waits -> commits issued)     median {np.median(r[5][i] - r[4][i]):7.0f} clk")
    t = np.arange(16, nt)
    print(f"  tile: last commit issued -> epilogue sees it   median {np.median(r[6][t] - r[5][t * nkc + nkc - 1]):7.0f} clk")
    print(f"  tile: epilogue duration (warp 8)               median {np.median(r[7][t] - r[6][t]):7.0f} clk")
    lead = [int(np.searchsorted(r[0][:512], x)) - k for k, x in zip(i, r[4][i])]
    print(f"  producer lead over MMA (chunks issued ahead)   median {np.median(lead):5.1f}")
    print("  chunks 200..211, clk since first issue (issue, sees X, slot, arrive, MMA go, commits):")
    for k in range(200, 212):
        print("   ", k, " ".join(f"{int(r[role][k] - t0):8d}" for role in (0, 1, 2, 3, 4, 5)))

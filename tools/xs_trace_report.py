"""Timeline report of a GCNB_XS_TRACE file (clock64 stamps of CTA 0 of the streaming X·W kernel).
    python tools/xs_trace_report.py <trace file>
Roles: 0 producer issued chunk i, 1 MMA warp at the top of chunk i's iteration, 2 MMA warp past the CONV wait,
3 last converter warp arrived on CONV, 4 MMA warp past tcgen05.fence, 5 MMA warp after the chunk's commits,
6 first epilogue warp saw tile t complete (TFULL), 7 first epilogue warp finished tile t."""
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
    for role, name in [(0, "producer issue"), (3, "conv arrive (last warp)"), (4, "MMA go"), (5, "MMA commits issued")]:
        d = np.diff(r[role][steady])
        print(f"  {name:24s} period/chunk: median {np.median(d):7.0f}  mean {d.mean():7.0f}  p90 {np.percentile(d, 90):7.0f} clk")
    nt = min(112, 448 // nkc)
    for role, name in [(6, "epilogue sees tile"), (7, "epilogue done")]:
        d = np.diff(r[role][16:nt])
        print(f"  {name:24s} period/tile : median {np.median(d):7.0f}  mean {d.mean():7.0f} clk  ({nkc} chunks per tile)")
    i = np.arange(64, 448)
    print(f"  X load + conversion (issue -> last converter arrive)  median {np.median(r[3][i] - r[0][i]):7.0f} clk")
    print(f"  MMA warp: loop top -> past CONV wait                   median {np.median(r[2][i] - r[1][i]):7.0f}  mean {np.mean(r[2][i] - r[1][i]):7.0f} clk")
    print(f"  MMA warp: CONV wait -> past fence                      median {np.median(r[4][i] - r[2][i]):7.0f}  mean {np.mean(r[4][i] - r[2][i]):7.0f} clk")
    print(f"  MMA warp: issue (fence -> commits issued)              median {np.median(r[5][i] - r[4][i]):7.0f}  mean {np.mean(r[5][i] - r[4][i]):7.0f} clk")
    print(f"  MMA warp: commits -> next loop top (incl. TEMPTY wait) median {np.median(r[1][i + 1] - r[5][i]):7.0f}  mean {np.mean(r[1][i + 1] - r[5][i]):7.0f} clk")
    print(f"  converters ahead of the MMA warp (arrive -> MMA go)    median {np.median(r[4][i] - r[3][i]):7.0f} clk")
    t = np.arange(16, nt)
    print(f"  tile: last commit issued -> epilogue sees it           median {np.median(r[6][t] - r[5][t * nkc + nkc - 1]):7.0f} clk")
    print(f"  tile: epilogue duration (first epilogue warp)          median {np.median(r[7][t] - r[6][t]):7.0f} clk")
    print("  chunks 200..211, clk since first issue (issue, conv arrive, MMA top, past CONV, past fence, commits):")
    for k in range(200, 212):
        print("   ", k, " ".join(f"{int(r[role][k] - t0):8d}" for role in (0, 3, 1, 2, 4, 5)))

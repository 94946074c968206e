"""Host time of the vertex orderings on the Reddit-shaped graph (generated on the CPU, seed 15 — the
graph the permutations cached in profiles/perm/ belong to), and a check that each ordering still
produces exactly the cached permutation.  Writes profiles/perm/reddit_host_seconds.json.

    python tools/host_order_times.py [deg rcm rabbit gorder]
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gcn_b200 import graph, synth  # noqa: E402

orders = sys.argv[1:] or ["deg", "rcm", "rabbit", "gorder"]
t0 = time.perf_counter()
rp, ci, va, n = synth.make_graph("reddit", device="cpu")
print(f"graph: n={n} nnz={int(rp[-1])} generated in {time.perf_counter() - t0:.1f} s", flush=True)
out_path = os.path.join(ROOT, "profiles", "perm", "reddit_host_seconds.json")
res = json.load(open(out_path)) if os.path.exists(out_path) else {}
if "single_thread_round1" not in res:      # keep the first measurement beside the new one
    res = {"single_thread_round1": {k: v for k, v in res.items() if isinstance(v, (int, float))}}
fn = {"deg": graph.renumber_deg, "rcm": graph.renumber_rcm, "rabbit": graph.renumber_rabbit,
      "gorder": graph.renumber_gorder}
now = res.setdefault("host_threads", {"threads": graph_threads if (graph_threads := os.environ.get("GCNB_HOST_THREADS")) else f"default ({os.cpu_count()} cpus)"})
for o in orders:
    t0 = time.perf_counter()
    _, _, _, vomp = fn[o](rp, ci, va)
    dt = round(time.perf_counter() - t0, 1)
    cached = np.load(os.path.join(ROOT, "profiles", "perm", f"reddit_{o}_vomp.npy"))
    same = bool(np.array_equal(vomp, cached))
    now[o] = dt
    now[o + "_equals_cached_permutation"] = same
    print(f"{o}: {dt} s, equals cached permutation: {same}", flush=True)
    json.dump(res, open(out_path, "w"), indent=1)

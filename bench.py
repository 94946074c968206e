#!/usr/bin/env python
"""Benchmark of the GCN aggregation SpMM hot path (BASELINE.json metric: SpMM GFLOP/s and HBM GB/s
against the roofline) on a synthetic graph of a BASELINE shape.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload products|arxiv|reddit|rmat]
                    [--k WIDTH] [--impl ours|reference]

A "step" is one pass of the hot path over the whole graph: C = Â·B (fp32), Â resident in HBM in
the plan's layout.  Default workload: the ogbn-products-shaped graph at feature width 100 — the
configuration BASELINE.json's target ("≥60 % of HBM roofline on ogbn-products-shaped graphs at 1
GPU") is quoted on and the one its 1/2/4/8-GPU row-partitioned runs use.  B (≈0.98 GB) is far
larger than L2, so every timed iteration streams from HBM; no explicit flush is needed.

N > 1 (launched by torchrun, one rank per GPU): the same graph, row-partitioned by non-zeros; the
remote rows of B arrive by our halo pull over NVLink peer memory (default, --exchange pull) or by an
NCCL all-gather of B (--exchange allgather), overlapped with the local-column block
(gcn_b200/dist.py) — strong scaling, value = total flops ÷ max-over-ranks device time.  Every N > 1
line carries an untimed `parity_check`: consecutive steps with fresh operands under the real
stream-ordered rank barrier, sampled rows against an fp64 host evaluation; the run exits non-zero
above the tolerance.

The default line also carries `epoch` (the second half of BASELINE.json's metric: one 2-layer GCN
training epoch on the same graph) and, at N = 1, `gpu_comparators` (cuSPARSE CSR ALG3 as the
reference's cuspmm.cu calls it, torch.spmm on CUDA as pygcn/gcn5.py:90 calls it, and with
--comparators all the reference's own flexspmm kernel on its corrected tiles) on the same inputs.

--impl reference: the reference's CPU implementation of the path — torch.spmm(adj, B) on COO
tensors (pygcn/gcn1.py:53) — on the box's host cores, on a bounded row-slice sample.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "spmm_gflops"
UNIT = "GFLOP/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="products", choices=["products", "arxiv", "reddit", "rmat", "cora"])
    ap.add_argument("--k", type=int, default=0, help="feature width (default: the shape's)")
    ap.add_argument("--scale", type=int, default=22, help="rmat: log2 vertices")
    ap.add_argument("--edge-factor", type=int, default=16)
    ap.add_argument("--dtype", default="f32", choices=["f32", "bf16"])
    ap.add_argument("--shuffle", action="store_true", help="random vertex relabelling (no id locality)")
    ap.add_argument("--pitch", type=int, default=0,
                    help="row pitch of B and C in elements (>= k; default k = contiguous rows)")
    ap.add_argument("--mode", default="spmm", choices=["spmm", "epoch"],
                    help="spmm: one C = A·B pass (headline); epoch: one 2-layer GCN training epoch")
    ap.add_argument("--order", default="none", choices=["none", "deg", "rcm", "gorder", "rabbit"],
                    help="host vertex reordering applied to the graph before the plan is built")
    ap.add_argument("--wire", default="f32", choices=["f32", "bf16"],
                    help="multi-GPU: dtype of the rows the ranks exchange (bf16 halves the NVLink bytes; the "
                         "result is then within the bf16 tolerance, 1e-2)")
    ap.add_argument("--exchange", default="pull", choices=["pull", "allgather"],
                    help="N>1: how remote feature rows arrive — our P2P halo pull over NVLink peer memory, "
                         "or an NCCL all-gather of the whole operand")
    ap.add_argument("--assoc", default="auto", choices=["auto", "reference"],
                    help="epoch mode: 'reference' = gc1 Â·(X·W1), gc2 (Â·H)·W2 as pygcn/gcn5.py builds them; 'auto' = "
                         "the same layers evaluated at the narrower SpMM width with Â·X cached (gcn_b200/model.py)")
    ap.add_argument("--perm-file", default="",
                    help="apply a cached vertex permutation (.npy int32 vomp[new] = old, e.g. a Gorder result "
                         "computed once on the host) instead of computing --order in this run")
    ap.add_argument("--cuda-graph", action="store_true",
                    help="epoch mode, 1 GPU: replay the training step from a CUDA graph (model.GCN.graphed_step)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-threads", type=int, default=0,
                    help="host threads of the reference arm / cpu_baseline (default: every core this process may "
                         "run on, also under torchrun, which exports OMP_NUM_THREADS=1)")
    ap.add_argument("--comparators", default="fast", choices=["none", "fast", "all"],
                    help="N=1 spmm mode: other GPU implementations timed on the same inputs (bench leg only): "
                         "fast = cuSPARSE CSR ALG3 + torch.spmm CUDA; all = + the reference's flexspmm.so on its tiles")
    ap.add_argument("--no-epoch", action="store_true", help="spmm mode: skip the `epoch` object of the line")
    ap.add_argument("--k-sweep", default="",
                    help="N=1 spmm mode: comma-separated feature widths timed on ONE generated graph / plan (the R-MAT "
                         "sweep of BASELINE config 5); prints one short JSON line per width instead of the bench line")
    ap.add_argument("--no-parity", action="store_true", help="N>1: skip the untimed parity_check")
    return ap.parse_args()


def host_threads(args):
    """Thread count of the CPU legs: every core the process may run on, set explicitly so that the N = 1
    and the torchrun (OMP_NUM_THREADS=1) arms time the reference on the same number of threads."""
    try:
        avail = len(os.sched_getaffinity(0))
    except AttributeError:
        avail = os.cpu_count() or 1
    n = args.cpu_threads if getattr(args, "cpu_threads", 0) > 0 else avail
    torch.set_num_threads(n)
    return torch.get_num_threads()


def kernel_launches_per_spmm(info, k, dtype):
    """Kernels one gcnb_spmm call launches (spmm.cu: launch_vec), counted from the plan: every width
    runs a stream-class kernel (packed for narrow rows) + one fixed-order reduction if a row was cut."""
    packed = int(os.environ.get("GCNB_PACKED", "8")) >= 2 and info["nnz"] >= 24 * info["m"]
    if packed or k > 64:
        return 1 + (info["stream_rows_split"] > 0)
    return 1 + (info["block_items"] > 0) + (info["rows_split"] > 0)


DEFAULT_K = {"products": 100, "arxiv": 256, "reddit": 128, "rmat": 128, "cora": 16}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (B200_PROFILING.md's clocks line), read
    in-process through NVML every 10 ms.  (A first version spawned `nvidia-smi` every 100 ms from every
    rank: with 8 ranks those processes took the driver lock in the middle of the timed steps and showed
    up as multi-millisecond outlier steps.)  Falls back to one nvidia-smi query at the end."""
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index=0, enabled=True):
        self.index, self.enabled = index, enabled
        self.sm, self.mx, self.bits, self.stop = [], None, 0, threading.Event()
        self.nv = self.h = None
        self.thread = threading.Thread(target=self._run, daemon=True)
        if enabled:
            try:
                import pynvml
                pynvml.nvmlInit()
                # NVML indexes physical devices; honour CUDA_VISIBLE_DEVICES when it lists indices
                vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
                ids = [int(x) for x in vis.split(",") if x.strip().isdigit()]
                phys = ids[index] if index < len(ids) else index
                self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
                self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
                self.nv = pynvml
            except Exception:
                self.nv = None

    def _sample(self):
        nv = self.nv
        try:
            self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
            self.bits |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
        except Exception:
            pass

    def _run(self):
        while not self.stop.is_set():
            self._sample()
            self.stop.wait(0.01)

    def __enter__(self):
        if self.enabled and self.nv is not None:
            self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        if self.thread.is_alive():
            self.thread.join(timeout=10)

    def _smi_once(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                                  str(self.index)], capture_output=True, text=True, timeout=10).stdout
            p = [x.strip() for x in out.strip().split(",")]
            return {"sm_mhz": float(p[0]), "sm_max_mhz": float(p[1]),
                    "reasons": [n for i, n in enumerate(self.NAMES) if p[2 + i].lower().startswith("active")],
                    "samples": 1, "source": "nvidia-smi after the timed region (NVML unavailable)"}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}

    def summary(self):
        if not self.enabled:
            return None
        if self.nv is None or not self.sm:
            return self._smi_once()
        nv = self.nv
        masks = [nv.nvmlClocksEventReasonHwSlowdown, nv.nvmlClocksEventReasonHwThermalSlowdown,
                 nv.nvmlClocksEventReasonSwThermalSlowdown, nv.nvmlClocksEventReasonSwPowerCap]
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.mx,
                "reasons": [n for n, m in zip(self.NAMES, masks) if self.bits & m],
                "samples": len(sm), "source": "NVML, 10 ms period, during the timed region"}


def make_workload(args, device):
    from gcn_b200 import synth
    k = args.k or DEFAULT_K[args.workload]
    if args.workload == "rmat":
        rp, ci, va, n = synth.make_graph("rmat", device=device, scale=args.scale,
                                         edge_factor=args.edge_factor, shuffle=args.shuffle)
        name = f"rmat-s{args.scale}-ef{args.edge_factor}"
    else:
        # a cached permutation belongs to ONE graph: torch's CPU and CUDA generators differ, so with
        # --perm-file the graph is generated on the CPU like the run that produced the file did
        gen_dev = "cpu" if args.perm_file else device
        rp, ci, va, n = synth.make_graph(args.workload, device=gen_dev, shuffle=args.shuffle)
        name = f"{args.workload}-shaped"
    args.reorder_s = None
    if args.perm_file:
        from gcn_b200 import graph
        vomp = np.load(args.perm_file).astype(np.int32)
        if sorted(vomp.tolist()) != list(range(n)):
            raise SystemExit(f"{args.perm_file} is not a permutation of {n} vertices")
        t0 = time.perf_counter()
        rp, ci, va = graph.perm_apply(rp, ci, va, vomp)
        args.reorder_s = round(time.perf_counter() - t0, 2)
        args.order = f"cached {os.path.basename(args.perm_file)}; applying it took"
    elif args.order != "none":
        # host preprocessing of the path (renumber.cu / order_*.cu equivalents, bit-exact): timed,
        # reported, outside the SpMM step
        from gcn_b200 import graph
        fn = {"deg": graph.renumber_deg, "rcm": graph.renumber_rcm, "gorder": graph.renumber_gorder,
              "rabbit": graph.renumber_rabbit}[args.order]
        t0 = time.perf_counter()
        rp, ci, va, _vomp = fn(rp, ci, va)
        args.reorder_s = round(time.perf_counter() - t0, 2)
    return name, rp, ci, va, n, k


def workload_string(name, k):
    return f"{name} SpMM C=A·B fp32, k={k}"


# ---- CPU baseline: the reference's path on the host cores ------------------------------------------
def cpu_reference_sample(rp, ci, va, n, k, budget_s=10.0, reps=1, args=None):
    """torch.spmm(adj, B) on CPU (pygcn/gcn1.py:53) over a bounded row slice of the same Â.
    Returns (GFLOP/s, cores, sample description, seconds per call)."""
    from oracle import spmm_oracle as SO
    if args is not None:
        host_threads(args)
    nnz = int(rp[-1])
    # slice sized for roughly budget_s at a guessed 4 GFLOP/s; at least 1/64 of the rows
    want_nnz = min(nnz, max(nnz // 64, int(budget_s * 2e9 / (2 * k))))
    hi = int(np.searchsorted(rp, want_nnz, side="left"))
    hi = max(1, min(n, hi))
    z = int(rp[hi])
    adj = SO.torch_coo(rp[: hi + 1], ci[:z], va[:z], (hi, n))
    B = torch.randn(n, k, generator=torch.Generator().manual_seed(15))
    SO.torch_spmm_cpu(adj, B)  # warm-up (thread pool, page faults)
    best = float("inf")
    for _ in range(reps):
        t0 = time.perf_counter()
        SO.torch_spmm_cpu(adj, B)
        best = min(best, time.perf_counter() - t0)
    gflops = 2.0 * z * k / best / 1e9
    sample = f"torch.spmm COO fp32 on rows [0,{hi}) of {n} ({z} of {nnz} nnz), k={k}, best of {reps}"
    return gflops, torch.get_num_threads(), sample, best


def run_reference_epoch(args):
    """--impl reference --mode epoch: the reference's 2-layer GCN epoch on the host cores (torch CPU ops,
    pygcn/gcn1.py:62-118, :176-203, restated in oracle/spmm_oracle.py gcn_epoch_cpu), whole graph —
    only for the shapes where that takes seconds (Cora- and arxiv-shaped)."""
    host_threads(args)
    name, rp, ci, va, n, _ = make_workload(args, "cuda" if torch.cuda.is_available() else "cpu")
    nnz = int(rp[-1])
    nfeat, nhid, nclass = EPOCH_SHAPES[args.workload]
    base = {"impl": "reference", "metric": "gcn_epoch_ms", "unit": "ms", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{name} 2-layer GCN {nfeat}->{nhid}->{nclass} training epoch "
                                   "(forward, NLL on 5 % of the vertices, backward, Adam); pygcn/gcn1.py:62-203",
                       "n": n, "nnz": nnz}}
    if nnz > 5_000_000:
        base["unavailable"] = "CPU epoch only for the Cora- and arxiv-shaped graphs (bounded run time); use --mode spmm"
        print(json.dumps(base))
        return
    from oracle import spmm_oracle as SO
    adj = SO.torch_coo(rp, ci, va, (n, n))
    g = torch.Generator().manual_seed(15)
    x = torch.randn(n, nfeat, generator=g)
    labels = torch.randint(0, nclass, (n,), generator=g)
    idx_train = torch.arange(n // 20)
    params = []
    for shape in ((nfeat, nhid), (nhid,), (nhid, nclass), (nclass,)):
        stdv = 1.0 / (shape[-1] ** 0.5)     # reset_parameters, gcn1.py:33-38
        params.append(((torch.rand(shape, generator=g) * 2 - 1) * stdv).requires_grad_(True))
    opt = torch.optim.Adam(params, lr=0.01, weight_decay=5e-4)
    for _ in range(args.warmup):
        SO.gcn_epoch_cpu(adj, x, labels, idx_train, params, opt)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        SO.gcn_epoch_cpu(adj, x, labels, idx_train, params, opt)
    ms = (time.perf_counter() - t0) / args.steps * 1e3
    base.update({"value": ms, "ms_per_step": ms,
                 "cpu_baseline": {"value": ms, "unit": "ms", "cores": torch.get_num_threads(), "kind": "port",
                                  "sample": "whole graph, every step a full epoch"},
                 "e2e": {"value": ms, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    print(json.dumps(base))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.mode == "epoch":
        run_reference_epoch(args)
        return
    host_threads(args)
    name, rp, ci, va, n, k = make_workload(args, "cuda" if torch.cuda.is_available() else "cpu")
    from oracle import spmm_oracle as SO
    nnz = int(rp[-1])
    want_nnz = min(nnz, max(nnz // 64, int(1.0 * 4e9 / (2 * k))))   # ≈1 s per step
    hi = max(1, min(n, int(np.searchsorted(rp, want_nnz, side="left"))))
    z = int(rp[hi])
    adj = SO.torch_coo(rp[: hi + 1], ci[:z], va[:z], (hi, n))
    B = torch.randn(n, k, generator=torch.Generator().manual_seed(15))
    for _ in range(args.warmup):
        SO.torch_spmm_cpu(adj, B)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        SO.torch_spmm_cpu(adj, B)
    dt = (time.perf_counter() - t0) / args.steps
    val = 2.0 * z * k / dt / 1e9
    sample = f"torch.spmm COO fp32 (pygcn/gcn1.py:53) on rows [0,{hi}) of {n} ({z} of {nnz} nnz), k={k}"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_string(name, k), "n": n, "nnz": nnz, "k": k},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": torch.get_num_threads(),
                         "kind": "reference", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ---- GPU comparators (bench leg only; nothing here is on the product path) ----------------------------
def _event_ms(fn, reps=5, warm=2):
    """Device time of fn() by CUDA events on the current (= legacy default) stream.  A ~5 ms spin kernel is
    queued first so that host-side stalls inside fn (the reference's per-call cudaGetDeviceProperties /
    cusparseCreate) happen while the GPU is still busy and do not show up between the two events."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(10_000_000)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        out.append(a.elapsed_time(b))
    out.sort()
    return out[len(out) // 2]


def gpu_comparators(args, rp, ci, va, n, k, B, C_ours, device, ours_ms):
    """Other GPU implementations of C = Â·B on the SAME device-resident inputs:
      cusparse_csr_alg3  cusparseSpMM(CSR, row-major dense, CUSPARSE_SPMM_CSR_ALG3) — the call of
                         /root/reference/cuspmm.cu:57-61, per-call setup hoisted (oracle/cusparse_cmp.cu)
      torch_spmm_cuda    torch.spmm(adj_coo, B) on CUDA — pygcn/gcn5.py:90, the reference's GPU layer
      ref_flexspmm       (--comparators all) the reference's own flexspmm.so (unmodified flexspmm.cu built
                         for sm_100a) on its corrected tile format, kernel time only
    Each entry: median ms of 5, GFLOP/s, max relative deviation of its result from ours."""
    import ctypes
    res = {}
    flops = 2.0 * int(rp[-1]) * k
    if args.dtype != "f32" or B.stride(0) != k:
        return {"skipped": "comparators run on contiguous fp32 operands only"}

    def entry(ms, C, note):
        err = float((C - C_ours).abs().max() / C_ours.abs().max())
        return {"ms": ms, "gflops": flops / (ms * 1e-3) / 1e9, "ours_over_this": ms / ours_ms,
                "max_rel_dev_from_ours": err, "note": note}

    d_rp = torch.from_numpy(np.ascontiguousarray(rp, dtype=np.int32)).to(device)
    d_ci = torch.from_numpy(np.ascontiguousarray(ci, dtype=np.int32)).to(device)
    d_va = torch.from_numpy(np.ascontiguousarray(va, dtype=np.float32)).to(device)
    Bc = B.contiguous()
    C = torch.empty(n, k, device=device)
    # -- cuSPARSE ALG3 (and ALG2, the one torch picks) with a persistent handle
    lib_path = os.path.join(ROOT, "oracle", "libcusparse_cmp.so")
    if os.path.exists(lib_path):
        lib = ctypes.CDLL(lib_path)
        lib.cmp_cusparse_create.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_longlong] * 3 + [ctypes.c_void_p] * 2 + \
            [ctypes.c_int, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        lib.cmp_cusparse_spmm.argtypes = [ctypes.c_void_p]
        lib.cmp_cusparse_destroy.argtypes = [ctypes.c_void_p]
        lib.cmp_cusparse_buffer_bytes.argtypes = [ctypes.c_void_p]
        lib.cmp_cusparse_buffer_bytes.restype = ctypes.c_longlong
        for alg, nm in ((0, "cusparse_csr_alg3"), (1, "cusparse_csr_alg2")):
            h = ctypes.c_void_p()
            rc = lib.cmp_cusparse_create(d_rp.data_ptr(), d_ci.data_ptr(), d_va.data_ptr(), n, n, int(rp[-1]),
                                         Bc.data_ptr(), C.data_ptr(), k, k, k, alg,
                                         torch.cuda.current_stream(device).cuda_stream, ctypes.byref(h))
            if rc != 0:
                res[nm] = {"error": f"cmp_cusparse_create rc={rc}"}
                continue
            ms = _event_ms(lambda: lib.cmp_cusparse_spmm(h))
            res[nm] = entry(ms, C, f"cusparseSpMM CSR x row-major fp32, work buffer {lib.cmp_cusparse_buffer_bytes(h)} B, "
                                   "handle/descriptors created once"
                                   + ("; the algorithm /root/reference/cuspmm.cu:57-61 names" if alg == 0 else ""))
            lib.cmp_cusparse_destroy(h)
    else:
        res["cusparse_csr_alg3"] = {"error": "oracle/libcusparse_cmp.so not built"}
    # -- torch.spmm on CUDA, COO adjacency as the reference builds it (utils.py:243-250)
    try:
        row = torch.repeat_interleave(torch.arange(n, device=device), (d_rp[1:] - d_rp[:-1]).long())
        adj = torch.sparse_coo_tensor(torch.stack((row, d_ci.long())), d_va, (n, n)).coalesce()
        del row
        holder = {}

        def f():
            holder["C"] = torch.spmm(adj, Bc)
        ms = _event_ms(f)
        res["torch_spmm_cuda"] = entry(ms, holder["C"], "torch.spmm(sparse COO fp32, dense) on CUDA, pygcn/gcn5.py:90")
        del adj, holder
    except Exception as e:   # noqa: BLE001 — a comparator that cannot run is reported, not fatal
        res["torch_spmm_cuda"] = {"error": repr(e)[:200]}
    # -- the reference's own kernel on its own (corrected) tiles
    if args.comparators == "all":
        from gcn_b200 import graph
        from oracle import ref_bindings as R
        if not R.have("flexspmm.so"):
            res["ref_flexspmm"] = {"error": "oracle/_ref/flexspmm.so not shipped"}
        elif n * k >= 2 ** 31 or n >= 2 ** 24:
            res["ref_flexspmm"] = {"error": "reference limits: int32 row*k addressing (flexspmm.cu:67), float column ids (tile.cu:67)"}
        else:
            t0 = time.perf_counter()
            n_sm = torch.cuda.get_device_properties(device).multi_processor_count
            t = graph.csr2tile(rp, ci, va, n_sm=n_sm)     # bit-exact with the corrected reference builder (tests)
            tile_s = time.perf_counter() - t0
            lib = ctypes.CDLL(os.path.join(R.REF_DIR, "flexspmm.so"))
            d = {key: torch.from_numpy(np.ascontiguousarray(t[key])).to(device)
                 for key in ("seg_rowPtr", "segNzCV", "segVoMap", "grouped_tailSeg", "next_seg")}
            Cf = torch.zeros(n, k, device=device)
            nxt = d["next_seg"].clone()
            vp = ctypes.c_void_p

            def f():
                Cf.zero_()                      # the reference allocates zeros and copies next_seg per call (gcn6.py:37-38)
                nxt.copy_(d["next_seg"])

            def g():
                lib.flexspmm(vp(d["seg_rowPtr"].data_ptr()), vp(d["segNzCV"].data_ptr()), vp(d["segVoMap"].data_ptr()),
                             vp(d["grouped_tailSeg"].data_ptr()), vp(nxt.data_ptr()), n, n, k, int(t["n_segs"]),
                             vp(Bc.data_ptr()), vp(Cf.data_ptr()))
            times = []
            for i in range(5):
                f()
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda._sleep(10_000_000)
                a.record()
                g()
                b.record()
                torch.cuda.synchronize()
                if i >= 2:
                    times.append(a.elapsed_time(b))
            times.sort()
            res["ref_flexspmm"] = entry(times[len(times) // 2], Cf,
                                        f"flexspmm.so (reference kernel, sm_100a build) on {int(t['n_segs'])} corrected tile-segments; "
                                        f"kernel only (zeroing C and the next_seg copy outside the events); host tile build {tile_s:.1f} s")
    return res


# ---- our arm -------------------------------------------------------------------------------------
def time_steps(fn, steps, warmup, device, barrier=None, align=None):
    """W untimed + K timed calls; per-step CUDA events on the launching (current) stream.
    `align` (N > 1): a stream-ordered rendezvous of the ranks enqueued right before the first event, so
    that the timed region starts on every GPU when the last rank arrives (a host-side barrier alone
    leaves the ranks up to a millisecond apart, which the first timed step then absorbs)."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(device)
    if barrier:
        barrier()
    torch.cuda.synchronize(device)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    if align:
        align()
    t0.record()
    for a, b in ev:
        a.record()
        fn()
        b.record()
    t1.record()
    torch.cuda.synchronize(device)
    if barrier:
        barrier()
    raw = [a.elapsed_time(b) for a, b in ev]
    time_steps.last = [round(x, 3) for x in raw]
    per = sorted(raw)
    return t0.elapsed_time(t1) / steps, per[len(per) // 2], per[0]


def run_single(args):
    from gcn_b200 import ops, synth
    device = torch.device("cuda", 0)
    torch.cuda.set_device(device)
    name, rp, ci, va, n, k = make_workload(args, device)
    nnz = int(rp[-1])
    t0 = time.perf_counter()
    plan = ops.SpmmPlan(rp, ci, va, (n, n), device=device, assume_symmetric=False, need_transpose=False)
    plan_s = time.perf_counter() - t0
    info = plan.info()
    dt = torch.float32 if args.dtype == "f32" else torch.bfloat16
    if args.k_sweep:
        peak, _ = peaks()
        with ClockSampler(0) as clk:
            for kk in [int(x) for x in args.k_sweep.split(",") if x]:
                g = torch.Generator(device=device).manual_seed(15)
                B = torch.randn(n, kk, device=device, generator=g).to(dt)
                C = torch.empty(n, kk, device=device, dtype=torch.float32)
                ms_avg, ms_med, ms_min = time_steps(lambda: plan.spmm(B, out=C), args.steps, args.warmup, device)
                abytes, flops = synth.spmm_bytes_flops(n, n, nnz, kk, b_bytes=4 if args.dtype == "f32" else 2)
                # linearity + checksum-of-checksums self-check at full size (untimed)
                colsum = torch.from_numpy(np.bincount(ci, weights=va.astype(np.float64), minlength=n)).to(device)
                lhs = C.double().sum(0)
                rhs = colsum @ B.double()
                chk = float((lhs - rhs).abs().max() / rhs.abs().max().clamp_min(1e-30))
                print(json.dumps({"metric": METRIC, "value": flops / (ms_avg * 1e-3) / 1e9, "unit": UNIT, "n_gpus": 1,
                                  "ms_per_step": ms_avg, "ms_median": ms_med, "ms_min": ms_min, "steps": args.steps,
                                  "dtype": args.dtype,
                                  "config": {"workload": workload_string(name, kk), "n": n, "nnz": nnz, "k": kk,
                                             "max_row_nnz": info["max_row_nnz"], "plan_build_s": round(plan_s, 2)},
                                  "roofline": {"bound": "hbm", "achieved": abytes / (ms_avg * 1e-3) / 1e9, "peak": peak,
                                               "unit": "GB/s", "frac": abytes / (ms_avg * 1e-3) / 1e9 / peak,
                                               "algorithmic_bytes_per_launch": abytes},
                                  "checksum_rel_err": chk}), flush=True)
                del B, C
                torch.cuda.empty_cache()
        return
    g = torch.Generator(device=device).manual_seed(15)
    pitch = max(args.pitch, k)
    B = torch.randn(n, pitch, device=device, generator=g).to(dt)[:, :k]
    C = torch.empty(n, pitch, device=device, dtype=torch.float32)[:, :k]
    # k >= 97 (more than 32 four-float vectors per row... see spmm.cu launch_vec): stream class = one
    # launch + the split-row reduction; narrower rows: lane-group rows kernel + block class + reduction
    launches_per_step = kernel_launches_per_spmm(info, k, args.dtype)

    with ClockSampler(0) as clk:
        ms_avg, ms_med, ms_min = time_steps(lambda: plan.spmm(B, out=C), args.steps, args.warmup, device)
    step_list = list(time_steps.last)
    b_bytes = 4 if args.dtype == "f32" else 2
    abytes, flops = synth.spmm_bytes_flops(n, n, nnz, k, b_bytes=b_bytes)
    value = flops / (ms_avg * 1e-3) / 1e9
    peak, peak_src = peaks()
    achieved = abytes / (ms_avg * 1e-3) / 1e9
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_avg, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": workload_string(name, k),
                   "n": n, "nnz": nnz, "k": k, "max_row_nnz": info["max_row_nnz"],
                   "order": ("random relabelling" if args.shuffle else "generator order")
                            + ("" if args.order == "none" else f" -> {args.order} ({args.reorder_s} s on the host)"),
                   "row_pitch_elems": pitch,
                   "l2": f"operands {abytes / 1e9:.2f} GB >> 126 MB L2; no flush needed",
                   "plan_build_s": round(plan_s, 2)},
        "gpu_launches": launches_per_step * args.steps,
        "ms_median": ms_med, "ms_min": ms_min, "ms_steps": step_list,
        "hbm_gbs_algorithmic": achieved, "pct_of_8TBs": 100 * achieved / 8000.0,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": abytes,
                     "note": f"whole-step device time of the {launches_per_step} launch(es) of one gcnb_spmm call (class kernel"
                             + (" + fixed-order split-row reduce" if launches_per_step > 1 else "")
                             + ") used as the launch duration; ncu launch lists under profiles/ give the shares"},
        "clocks": clk.summary(),
    }
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_file):
        try:
            with open(traffic_file) as f:
                tr = json.load(f)
            key = f"{args.workload}-k{k}-{args.dtype}"
            if key in tr:
                # dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel from the committed
                # `ncu --set full` capture named in traffic_source — NOT measured in this run
                out["roofline"]["traffic"] = tr[key]["dram_bytes_per_launch"]
                out["roofline"]["traffic_source"] = tr[key].get("source")
                out["roofline"]["traffic_capture"] = tr[key].get("capture", "round 1")
        except Exception:
            pass

    if not args.no_e2e and args.dtype == "f32":
        # end to end through the C ABI with HOST buffers: H2D of B, SpMM, D2H of C inside the call
        Bh = torch.randn(n, k, generator=torch.Generator().manual_seed(15)).pin_memory()
        Ch = torch.empty(n, k).pin_memory()
        Bn, Cn = Bh.numpy(), Ch.numpy()
        reps = max(3, min(args.steps, 10))
        plan.spmm_host(Bn, out=Cn)
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        for _ in range(reps):
            plan.spmm_host(Bn, out=Cn)     # synchronises before returning
        e2e_s = (time.perf_counter() - t0) / reps
        out["e2e"] = {"value": flops / e2e_s / 1e9, "unit": UNIT, "h2d_bytes_per_step": n * k * 4,
                      "d2h_bytes_per_step": n * k * 4, "ms_per_step": e2e_s * 1e3, "steps": reps,
                      "api": "gcnb_spmm_host (C ABI, pinned host buffers)"}
        del Bh, Ch
    if args.comparators != "none":
        plan.spmm(B, out=C)
        try:
            out["gpu_comparators"] = gpu_comparators(args, rp, ci, va, n, k, B, C, device, ms_med)
        except Exception as e:   # noqa: BLE001
            out["gpu_comparators"] = {"error": repr(e)[:300]}
    del plan, B, C
    torch.cuda.empty_cache()
    if not args.no_epoch and args.workload in EPOCH_SHAPES:
        try:
            out["epoch"] = epoch_object_single(args, rp, ci, va, n, device)
        except Exception as e:   # noqa: BLE001
            out["epoch"] = {"error": repr(e)[:300]}
        torch.cuda.empty_cache()
    if not args.no_cpu_baseline:
        v, cores, sample, _ = cpu_reference_sample(rp, ci, va, n, k, args=args)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample}
    print(json.dumps(out))


EPOCH_SHAPES = {"cora": (1433, 16, 7), "arxiv": (128, 256, 40), "reddit": (602, 128, 41),
                "products": (100, 256, 47), "rmat": (128, 128, 16)}


def epoch_line(args, name, n, nnz, world, ms_avg, ms_med, ms_min, clocks, extra):
    nfeat, nhid, nclass = EPOCH_SHAPES[args.workload]
    out = {"metric": "gcn_epoch_ms", "value": ms_avg, "unit": "ms", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms_avg, "higher_is_better": False, "scaling": "strong",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": f"{name} 2-layer GCN {nfeat}->{nhid}->{nclass} training epoch "
                                  "(forward, NLL on 5 % of the vertices, backward, Adam); pygcn/gcn5.py:146-262",
                      "association": args.assoc, "cuda_graph": bool(getattr(args, "cuda_graph", False)),
                      "n": n, "nnz": nnz, "l2": "operands >> L2; no flush needed"},
           "ms_median": ms_med, "ms_min": ms_min, "clocks": clocks, "e2e": None}
    out.update(extra)
    return out


def epoch_object_single(args, rp, ci, va, n, device):
    """The second half of BASELINE.json's metric ("GCN epoch ms") on the SAME graph, for the default
    line: one 2-layer GCN training epoch (forward, NLL on 5 % of the vertices, backward, Adam),
    pygcn/gcn5.py:146-262 / gcn6.py:412-441, device-timed, plus end to end with the loss read back."""
    from gcn_b200 import model as gmodel
    from gcn_b200 import ops
    nfeat, nhid, nclass = EPOCH_SHAPES[args.workload]
    plan = ops.SpmmPlan(rp, ci, va, (n, n), device=device, assume_symmetric=True)
    g = torch.Generator(device=device).manual_seed(15)
    X = torch.randn(n, nfeat, device=device, generator=g)
    labels = torch.randint(0, nclass, (n,), device=device, generator=g)
    idx_train = torch.arange(n // 20, device=device)
    torch.manual_seed(15)
    net = gmodel.GCN(nfeat, nhid, nclass, dataset=args.workload, device=device, association=args.assoc)
    net.features, net.adj_norm, net.labels = X, plan, labels
    opt = torch.optim.Adam(net.parameters(), lr=net.lr, weight_decay=net.weight_decay)
    step = lambda: net._step(opt, labels, idx_train)   # noqa: E731
    steps = max(3, min(args.steps, 10))
    ms_avg, ms_med, ms_min = time_steps(step, steps, 3, device)
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    for _ in range(steps):
        loss = float(step()[1].item())
    e2e_ms = (time.perf_counter() - t0) / steps * 1e3
    return {"metric": "gcn_epoch_ms", "ms_per_epoch": ms_avg, "ms_median": ms_med, "ms_min": ms_min, "steps": steps,
            "warmup": 3, "shape": f"{nfeat}->{nhid}->{nclass}", "association": args.assoc, "loss": loss,
            "e2e_ms_per_epoch": e2e_ms,
            "e2e": "GCN._step + loss.item() every epoch; features, labels and the plan resident as in the reference's fit()"}


def epoch_object_multi(args, part, n, device, barrier):
    """Same as epoch_object_single on the row partition (gcn_b200.dist.PartitionedGCN); max over ranks."""
    import torch.distributed as dist
    from gcn_b200 import dist as gdist
    nfeat, nhid, nclass = EPOCH_SHAPES[args.workload]
    g = torch.Generator(device=device).manual_seed(15 + part.rank)
    X = torch.randn(part.m_local, nfeat, device=device, generator=g)
    labels = torch.randint(0, nclass, (part.m_local,), device=device, generator=g)
    n_train = n // 20
    idx_train = torch.arange(max(0, min(part.hi, n_train) - part.lo), device=device)
    net = gdist.PartitionedGCN(part, nfeat, nhid, nclass, association=args.assoc)
    steps = max(3, min(args.steps, 10))
    ms_avg, ms_med, ms_min = time_steps(lambda: net.step(X, labels, idx_train, n_train), steps, 3, device, barrier,
                                        align=part._rank_barrier)
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        loss = net.step(X, labels, idx_train, n_train).detach().clone()
        dist.all_reduce(loss, group=part.group)
        lv = float(loss.item())
    barrier()
    e2e_ms = (time.perf_counter() - t0) / steps * 1e3
    t = torch.tensor([ms_avg, ms_med, ms_min, e2e_ms], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_avg, ms_med, ms_min, e2e_ms = t.tolist()
    return {"metric": "gcn_epoch_ms", "ms_per_epoch": ms_avg, "ms_median": ms_med, "ms_min": ms_min, "steps": steps,
            "warmup": 3, "shape": f"{nfeat}->{nhid}->{nclass}", "association": args.assoc, "loss": lv,
            "e2e_ms_per_epoch": e2e_ms,
            "e2e": "PartitionedGCN.step + all-reduced loss read back every epoch; max over ranks"}


def run_epoch_single(args):
    from gcn_b200 import model as gmodel
    from gcn_b200 import ops
    device = torch.device("cuda", 0)
    torch.cuda.set_device(device)
    name, rp, ci, va, n, _ = make_workload(args, device)
    nnz = int(rp[-1])
    nfeat, nhid, nclass = EPOCH_SHAPES[args.workload]
    plan = ops.SpmmPlan(rp, ci, va, (n, n), device=device, assume_symmetric=True)
    g = torch.Generator(device=device).manual_seed(15)
    X = torch.randn(n, nfeat, device=device, generator=g)
    labels = torch.randint(0, nclass, (n,), device=device, generator=g)
    idx_train = torch.arange(n // 20, device=device)
    torch.manual_seed(15)
    net = gmodel.GCN(nfeat, nhid, nclass, dataset=args.workload, device=device, association=args.assoc)
    net.features, net.adj_norm, net.labels = X, plan, labels
    if args.cuda_graph:
        opt = torch.optim.Adam(net.parameters(), lr=net.lr, weight_decay=net.weight_decay, capturable=True)
        step = net.graphed_step(opt, labels, idx_train, warmup=3)
    else:
        opt = torch.optim.Adam(net.parameters(), lr=net.lr, weight_decay=net.weight_decay)
        step = lambda: net._step(opt, labels, idx_train)   # noqa: E731
    with ClockSampler(0) as clk:
        ms_avg, ms_med, ms_min = time_steps(step, args.steps, args.warmup, device)
    # end to end as the reference's training loop runs it (gcn6.py:416-441): features, labels and Â stay
    # resident (uploaded once by fit, gcn6.py:358), every epoch ends with the loss read back on the host
    reps = max(3, min(args.steps, 10))
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    for _ in range(reps):
        float(step()[1].item())
    e2e_ms = (time.perf_counter() - t0) / reps * 1e3
    print(json.dumps(epoch_line(args, name, n, nnz, 1, ms_avg, ms_med, ms_min, clk.summary(),
                                {"e2e": {"value": e2e_ms, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 4,
                                         "api": "GCN._step / graphed_step + loss.item() per epoch; inputs resident "
                                                "as in the reference's fit()"},
                                 "gpu_launches": None, "note": ("2 SpMMs (width nclass) + 6 tcgen05 GEMMs of ours per epoch, Â·X cached; "
                                                                if args.assoc == "auto" else "4 SpMMs + 6 tcgen05 GEMMs of ours per epoch; ")
                                 + "dropout / log_softmax / NLL / Adam are torch elementwise kernels"})))


def run_epoch_multi(args):
    import torch.distributed as dist
    from gcn_b200 import dist as gdist
    rank = int(os.environ["RANK"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    world = int(os.environ["WORLD_SIZE"])
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    dist.init_process_group("nccl", device_id=device)
    name, rp, ci, va, n, _ = make_workload(args, device)
    nnz = int(rp[-1])
    nfeat, nhid, nclass = EPOCH_SHAPES[args.workload]
    part = gdist.PartitionedSpMM(rp, ci, va, n, device=device, symmetric=True, exchange=args.exchange,
                                 wire_dtype=torch.bfloat16 if args.wire == "bf16" else None)
    g = torch.Generator(device=device).manual_seed(15 + rank)
    X = torch.randn(part.m_local, nfeat, device=device, generator=g)
    labels = torch.randint(0, nclass, (part.m_local,), device=device, generator=g)
    n_train = n // 20
    idx_train = torch.arange(max(0, min(part.hi, n_train) - part.lo), device=device)
    net = gdist.PartitionedGCN(part, nfeat, nhid, nclass, association=args.assoc)

    def barrier():
        dist.barrier(device_ids=[local])

    with ClockSampler(local) as clk:
        ms_avg, ms_med, ms_min = time_steps(lambda: net.step(X, labels, idx_train, n_train), args.steps,
                                            args.warmup, device, barrier, align=part._rank_barrier)
    t = torch.tensor([ms_avg, ms_med, ms_min], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    parity = None
    if not args.no_parity:   # the exchange this epoch rides on, checked at the hidden width
        parity = partition_parity_check(part, rp, ci, va, nhid, device, tol=1e-2 if args.wire == "bf16" else 1e-5)
    if rank == 0:
        ms_avg, ms_med, ms_min = t.tolist()
        print(json.dumps(epoch_line(args, name, n, nnz, world, ms_avg, ms_med, ms_min, clk.summary(),
                                    {"gpu_launches": None, "parity_check": parity,
                                     "parallelism": f"1-D row partition by nnz over {world} GPUs; "
                                     f"4 exchanges of the dense operand ({args.exchange}) + 1 gradient all-reduce per epoch"})))
    dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        raise SystemExit(f"bench.py: parity_check failed: {parity['max_rel_err']:.3e} > {parity['tol']}")


def partition_parity_check(part, rp, ci, va, k, device, tol, steps=6, rows=256):
    """Untimed self-check of the N-GPU path: `steps` consecutive partitioned SpMMs
    under the REAL stream-ordered rank barrier (gcn_b200.dist._rank_barrier), each with
    a fresh operand (so that a stale exchange slot cannot pass; both slots are reused from step 3 on), and
    `rows` sampled rows of this rank's result compared with an fp64 host evaluation of the same rows of
    Â·B.  Returns {"max_rel_err", "rows", "steps", "tol", "ok"}, the error being the max over ranks of
    max|C - C64| / max|C64|."""
    import torch.distributed as dist
    rank, world, stride, bounds = part.rank, part.world, part.stride, part.bounds
    lo = part.lo
    rng = np.random.default_rng(1234 + rank)
    worst = 0.0
    n_rows = 0
    # operands and, for the reference only, every rank's operand (a plain NCCL all-gather of the padded blocks),
    # all prepared BEFORE the steps, so that the steps themselves run back to back without a host synchronisation
    # between them: a rank may be a whole step ahead of a peer, which is when a stale or overwritten slot would show
    Bs, fulls = [], []
    for s in range(steps):
        g = torch.Generator(device=device).manual_seed(1000 * (s + 1) + rank)
        B = torch.randn(part.m_local, k, device=device, generator=g)
        padded = torch.zeros(stride, k, device=device)
        padded[: part.m_local] = B
        full = torch.empty(world * stride, k, device=device)
        dist.all_gather_into_tensor(full, padded, group=part.group)
        Bs.append(B)
        fulls.append(full)
    torch.cuda.synchronize(device)
    outs = []
    for s in range(steps):
        slot = part.next_operand(k)
        if slot is not None:        # fp32 pull mode: the producer writes into the free exchange slot
            slot.copy_(Bs[s])
            op = slot
        else:
            op = Bs[s]
        outs.append(part.forward(op))
    torch.cuda.synchronize(device)
    for s in range(steps):
        out, full = outs[s], fulls[s]
        pick = np.sort(rng.choice(part.m_local, size=min(rows, part.m_local), replace=False))
        got = out[torch.from_numpy(pick).to(device)].double().cpu().numpy()
        ref = np.zeros((len(pick), k))
        for i, r in enumerate(pick):
            z0, z1 = int(rp[lo + r]), int(rp[lo + r + 1])
            cols = ci[z0:z1].astype(np.int64)
            owner = np.searchsorted(bounds, cols, side="right") - 1
            pos = owner * stride + (cols - bounds[owner])
            brow = full[torch.from_numpy(pos).to(device)].double().cpu().numpy()
            ref[i] = va[z0:z1].astype(np.float64) @ brow
        worst = max(worst, float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30)))
        n_rows += len(pick)
    t = torch.tensor([worst], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=part.group)
    cnt = torch.tensor([n_rows], device=device, dtype=torch.float64)
    dist.all_reduce(cnt, op=dist.ReduceOp.SUM, group=part.group)
    err = float(t.item())
    return {"max_rel_err": err, "rows": int(cnt.item()), "steps": steps, "tol": tol, "ok": bool(err <= tol),
            "barrier": "gcn_b200.dist._rank_barrier (" + ("flags in peer memory, gcnb_peer_barrier"
                                                        if getattr(part, "_peer_flags", None) is not None
                                                        else "stream-ordered NCCL all-reduce")
                       + "), fresh operand every step, steps back to back without a host synchronisation",
            "reference": "fp64 host evaluation of the sampled rows of A·B"}


def run_multi(args):
    import torch.distributed as dist
    from gcn_b200 import dist as gdist
    from gcn_b200 import synth
    rank = int(os.environ["RANK"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    world = int(os.environ["WORLD_SIZE"])
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    dist.init_process_group("nccl", device_id=device)
    name, rp, ci, va, n, k = make_workload(args, device)   # same seed on every rank -> same graph
    nnz = int(rp[-1])
    wire = torch.bfloat16 if args.wire == "bf16" else None
    in_slot = args.exchange == "pull" and wire is None   # fp32 operands are produced inside the exchange slot
    part = gdist.PartitionedSpMM(rp, ci, va, n, device=device, symmetric=True, exchange=args.exchange,
                                 wire_dtype=wire)
    local_nnz = int(part.fwd["local_csr"][0][-1])
    remote_nnz = int(part.fwd["remote_csr"][0][-1])
    g = torch.Generator(device=device).manual_seed(15 + rank)
    B = torch.randn(part.m_local, k, device=device, generator=g)
    if in_slot:
        # the operand is resident where its producer would have written it: the exchange slot
        # (gcn_b200.dist.PartitionedGCN does that with the X·W product)
        slot = part.next_operand(k)
        slot.copy_(B)
        B = slot

    def barrier():
        dist.barrier(device_ids=[local])

    with ClockSampler(local) as clk:
        ms_avg, ms_med, ms_min = time_steps(lambda: part.forward(B), args.steps, args.warmup, device, barrier,
                                            align=part._rank_barrier)
    step_list = time_steps.last
    # the three phases of a step, each timed alone on the compute stream (untimed region of the bench):
    # what the overlap has to hide, and what it cannot
    blocks = part.fwd
    out_l = torch.empty((part.m_local, k), device=device)
    ph_local = time_steps(lambda: part._mul_local(blocks, B, out_l), 5, 2, device, barrier)[1]
    holder = {}

    def _ex():
        holder["full"] = part._exchange(blocks, B, part._prepare(B))
    ph_exch = time_steps(_ex, 5, 2, device, barrier)[1]
    ph_remote = time_steps(lambda: part._mul_remote_accumulate(blocks, holder["full"], out_l), 5, 2, device, barrier)[1]
    # end to end through the public API (gcn_b200.dist): this rank's rows of B from pinned host memory,
    # the partitioned SpMM, this rank's rows of C back to pinned host memory, every step
    e2e_ms, e2e_parts = 0.0, [0.0, 0.0, 0.0]
    if not args.no_e2e:
        Bh = torch.randn(part.m_local, k, generator=torch.Generator().manual_seed(15 + rank)).pin_memory()
        Ch = torch.empty(part.m_local, k).pin_memory()
        dst_ag = None if in_slot else torch.empty(part.m_local, k, device=device)

        def e2e_step():
            # pull mode: the upload lands in the free exchange slot, alternating like any producer does
            # (a peer may still be pulling the slot of the previous step)
            dst = part.next_operand(k) if dst_ag is None else dst_ag
            dst.copy_(Bh, non_blocking=True)
            Ch.copy_(part.forward(dst), non_blocking=True)
            torch.cuda.synchronize(device)
        e2e_step()
        reps = max(3, min(args.steps, 10))
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            e2e_step()
        barrier()
        e2e_ms = (time.perf_counter() - t0) / reps * 1e3
        # where the time goes (this rank, device events): host->device, partitioned SpMM, device->host
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        dst = part.next_operand(k) if dst_ag is None else dst_ag
        evs[0].record(); dst.copy_(Bh, non_blocking=True)
        evs[1].record(); out_e = part.forward(dst)
        evs[2].record(); Ch.copy_(out_e, non_blocking=True)
        evs[3].record(); torch.cuda.synchronize(device)
        e2e_parts = [evs[i].elapsed_time(evs[i + 1]) for i in range(3)]
    parity = None
    if not args.no_parity:
        parity = partition_parity_check(part, rp, ci, va, k, device, tol=1e-2 if args.wire == "bf16" else 1e-5)
    epoch = None
    if not args.no_epoch and args.workload in EPOCH_SHAPES:
        epoch = epoch_object_multi(args, part, n, device, barrier)
    halo_rows = int(blocks["halo_row"].numel()) if args.exchange == "pull" else (world - 1) * part.stride
    # OUR kernels per step on this rank (NCCL's own kernels in all-gather mode are not counted): the two blocks'
    # SpMM launches (+ their split-row reductions), in pull mode the halo pull and, between GPUs, the rendezvous kernel
    launches_per_step = (kernel_launches_per_spmm(blocks["local"].info(), k, "f32")
                         + kernel_launches_per_spmm(blocks["remote"].info(), k, args.wire))
    if args.exchange == "pull":
        launches_per_step += 1 + (1 if getattr(part, "_peer_flags", None) is not None else 0)
    t = torch.tensor([ms_avg, ms_med, ms_min, ph_local, ph_exch, ph_remote, halo_rows, e2e_ms], device=device,
                     dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    stats = torch.tensor([local_nnz, remote_nnz], device=device, dtype=torch.float64)
    dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    ms_avg, ms_med, ms_min, ph_local, ph_exch, ph_remote, halo_rows, e2e_ms = t.tolist()
    if rank == 0:
        abytes, flops = synth.spmm_bytes_flops(n, n, nnz, k)
        peak, peak_src = peaks()
        achieved = abytes / (ms_avg * 1e-3) / 1e9
        gathered = int(halo_rows) * k * (2 if args.wire == "bf16" else 4)
        print(json.dumps({
            "metric": METRIC, "value": flops / (ms_avg * 1e-3) / 1e9, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_avg, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None,
            "dtype": "f32" if args.wire == "f32" else "f32 (remote rows rounded to bf16 on the wire)",
            "data": "synthetic",
            "config": {"workload": f"{name} SpMM C=A·B, k={k}, 1-D row partition by nnz over {world} GPUs, "
                                   + ("halo rows pulled over NVLink peer memory (gcnb_pull_rows)"
                                      if args.exchange == "pull" else "NCCL all-gather of B")
                                   + " overlapped with the local-column block"
                                   + ("; the operand is resident in the rank's exchange slot when the timed region "
                                      "starts (where its producer, X·W, writes it: no publish copy in the step)"
                                      if in_slot else ""),
                       "n": n, "nnz": nnz, "k": k, "exchange": args.exchange, "wire": args.wire,
                       "local_nnz_frac": stats[0].item() / max(nnz, 1),
                       "exchange_bytes_received_per_rank_max": gathered,
                       "l2": "operands >> L2; no flush needed"},
            "ms_steps_rank0": step_list,
            "phase_ms_alone": {"local_block_spmm": ph_local, "exchange": ph_exch, "remote_block_spmm": ph_remote,
                               "note": "median of 5, each phase run alone, max over ranks; the step overlaps the first two"},
            "gpu_launches": launches_per_step * args.steps, "ms_median": ms_med, "ms_min": ms_min,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak * world, "unit": "GB/s",
                         "frac": achieved / (peak * world), "traffic": None, "peak_source": peak_src,
                         "note": "aggregate: algorithmic bytes of the whole SpMM over N x per-GPU HBM peak; "
                                 "the step also moves allgather_bytes over NVLink"},
            "clocks": clk.summary(),
            "parity_check": parity, "epoch": epoch,
            "e2e": None if args.no_e2e else {
                "value": flops / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": n * k * 4,
                "d2h_bytes_per_step": n * k * 4, "ms_per_step": e2e_ms,
                "rank0_ms_h2d_spmm_d2h": [round(x, 3) for x in e2e_parts],
                "api": "gcn_b200.dist.PartitionedSpMM.forward per rank, pinned host rows in and out; bytes are "
                       "the sum over ranks, time the max over ranks (wall clock between barriers)"},
        }))
    dist.destroy_process_group()
    if parity is not None and not parity["ok"]:
        raise SystemExit(f"bench.py: parity_check failed: {parity['max_rel_err']:.3e} > {parity['tol']}")


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the hot path is CUDA only (no CPU fallback); "
                         "use --impl reference for the CPU reference arm")
    if args.mode == "epoch":
        (run_epoch_multi if world > 1 else run_epoch_single)(args)
    elif world > 1:
        run_multi(args)
    else:
        run_single(args)


if __name__ == "__main__":
    main()

"""bench.py's one-line JSON contract, as far as it can be exercised without a GPU: the reference arm
(CPU torch.spmm, pygcn/gcn1.py:53) on the Cora-shaped config, and the refusal to run our arm on CPU."""
import json
import os
import subprocess
import sys

import pytest
import torch

from conftest import ROOT

BENCH = os.path.join(ROOT, "bench.py")


def _run(*args):
    env = dict(os.environ, OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    return subprocess.run([sys.executable, BENCH, *args], capture_output=True, text=True, timeout=600, env=env)


def test_reference_arm_line():
    r = _run("--impl", "reference", "--workload", "cora", "--steps", "2", "--warmup", "1", "--gpus", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1                      # ONE JSON line
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "spmm_gflops" and d["unit"] == "GFLOP/s"
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 1 and d["higher_is_better"] is True
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["vs_baseline"] is None and d["data"] == "synthetic"
    assert "cora-shaped" in d["config"]["workload"] and d["config"]["k"] == 16
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and cb["value"] == d["value"] and "torch.spmm" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_stay_silent():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, BENCH, "--impl", "reference", "--workload", "cora", "--gpus", "2"],
                       capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_our_arm_refuses_to_run_without_a_gpu():
    r = _run("--workload", "cora")
    assert r.returncode != 0 and "no CPU fallback" in (r.stderr + r.stdout)


def test_kernel_launch_count_follows_the_dispatch():
    sys.path.insert(0, ROOT)
    import bench
    dense = {"nnz": 5000, "m": 100, "stream_rows_split": 1, "block_items": 3, "rows_split": 1}
    short = {"nnz": 500, "m": 100, "stream_rows_split": 0, "block_items": 0, "rows_split": 0}
    assert bench.kernel_launches_per_spmm(dense, 100, "f32") == 2      # stream kernel + split-row reduce
    assert bench.kernel_launches_per_spmm(dense, 32, "f32") == 2       # packed stream kernel + reduce
    assert bench.kernel_launches_per_spmm(short, 32, "f32") == 1       # lane-group kernel only
    assert bench.kernel_launches_per_spmm(short, 256, "bf16") == 1


def test_reference_epoch_arm_line():
    r = _run("--impl", "reference", "--mode", "epoch", "--workload", "cora", "--steps", "2", "--warmup", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][0])
    assert d["impl"] == "reference" and d["metric"] == "gcn_epoch_ms" and d["higher_is_better"] is False
    assert d["value"] > 0 and d["cpu_baseline"]["kind"] == "port" and "1433->16->7" in d["config"]["workload"]
    # shapes whose CPU epoch would take minutes are declined, not sampled
    r = _run("--impl", "reference", "--mode", "epoch", "--workload", "rmat", "--scale", "18", "--edge-factor", "16")
    d = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][0])
    assert r.returncode == 0 and "unavailable" in d


def test_defaults_and_multi_gpu_switches(monkeypatch):
    """With no flags: N = 1 and a K/W that finish within minutes (W >= 3); the multi-GPU switches parse."""
    sys.path.insert(0, ROOT)
    import bench
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    a = bench.parse()
    assert a.gpus == 1 and a.warmup >= 3 and a.steps >= 1 and a.impl == "ours" and a.mode == "spmm"
    assert a.workload == "products" and a.exchange == "pull" and a.wire == "f32"
    monkeypatch.setattr(sys, "argv", ["bench.py", "--gpus", "8", "--exchange", "allgather", "--wire", "bf16",
                                      "--mode", "epoch"])
    a = bench.parse()
    assert (a.gpus, a.exchange, a.wire, a.mode) == (8, "allgather", "bf16", "epoch")

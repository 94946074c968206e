"""Where one step of the row-partitioned SpMM spends its time when the phases overlap: CUDA events on both streams
inside PartitionedSpMM._apply (barrier -> pull on the side stream, local block on the compute stream, then the remote
block), median over the timed steps, every rank's line gathered on rank 0.
  python -m torch.distributed.run --nproc-per-node N tools/dist_timeline.py [--k 100] [--steps 20]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--k", type=int, default=100)
ap.add_argument("--steps", type=int, default=20)
args = ap.parse_args()
rank = int(os.environ["RANK"]); local = int(os.environ.get("LOCAL_RANK", rank)); world = int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
from gcn_b200 import dist as gdist, synth

rp, ci, va, n = synth.make_graph("products", device=dev)
part = gdist.PartitionedSpMM(rp, ci, va, n, device=dev, symmetric=True, exchange="pull")
k = args.k
B = part.next_operand(k)
B.copy_(torch.randn(part.m_local, k, device=dev, generator=torch.Generator(device=dev).manual_seed(15 + rank)))
blocks = part.fwd
E = lambda: torch.cuda.Event(enable_timing=True)
rows = []


def step():
    cur = torch.cuda.current_stream(dev)
    ev = {name: E() for name in ("start", "bar0", "bar1", "pull1", "loc0", "loc1", "rem0", "rem1")}
    out = torch.empty((part.m_local, k), dtype=torch.float32, device=dev)
    ev["start"].record(cur)
    pre = part._prepare(B)
    part._side.wait_stream(cur)
    with torch.cuda.stream(part._side):
        ex, slot = pre
        ev["bar0"].record(part._side)
        part._rank_barrier()
        ev["bar1"].record(part._side)
        after = torch.cuda.Event(); after.record(part._side)
        full = part._pull(blocks, ex, slot, k)
        ev["pull1"].record(part._side)
    if part._local_after_barrier(blocks, B):
        cur.wait_event(after)
    ev["loc0"].record(cur)
    part._mul_local(blocks, B, out)
    ev["loc1"].record(cur)
    cur.wait_stream(part._side)
    ev["rem0"].record(cur)
    part._mul_remote_accumulate(blocks, full, out)
    ev["rem1"].record(cur)
    return ev


for _ in range(5):
    step()
torch.cuda.synchronize(dev)
dist.barrier(device_ids=[local])
part._rank_barrier()
evs = [step() for _ in range(args.steps)]
torch.cuda.synchronize(dev)
s = evs[0]["start"]
med = lambda f: float(np.median([f(e) for e in evs]))
line = {"rank": rank, "world": world, "k": k,
        "step_ms": med(lambda e: e["start"].elapsed_time(e["rem1"])),
        "barrier_ms": med(lambda e: e["bar0"].elapsed_time(e["bar1"])),
        "pull_ms": med(lambda e: e["bar1"].elapsed_time(e["pull1"])),
        "local_ms": med(lambda e: e["loc0"].elapsed_time(e["loc1"])),
        "start_to_pull_end_ms": med(lambda e: e["start"].elapsed_time(e["pull1"])),
        "start_to_local_end_ms": med(lambda e: e["start"].elapsed_time(e["loc1"])),
        "start_to_remote_start_ms": med(lambda e: e["start"].elapsed_time(e["rem0"])),
        "remote_ms": med(lambda e: e["rem0"].elapsed_time(e["rem1"])),
        "whole_ms": evs[0]["start"].elapsed_time(evs[-1]["rem1"]) / args.steps,
        "env": {k_: v for k_, v in os.environ.items() if k_.startswith("GCNB_")}}
out = [None] * world
dist.all_gather_object(out, line)
if rank == 0:
    for l in out:
        print(json.dumps(l))
part.close()
dist.destroy_process_group()

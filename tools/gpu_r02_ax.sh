#!/bin/bash
# round 2, multi-GPU batch ax: L2 eviction priority on the halo rows the pull writes (and on its peer loads): does the
# local block keep its L2 hit rate beside the pull?  (GCNB_PULL_HINT bit 0 = stores, bit 1 = loads)
tag=${1:-r02_ax}; N=${2:-4}; o=gpurun_out; mkdir -p $o
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
for e in "GCNB_PULL_HINT=0" "GCNB_PULL_HINT=1" "GCNB_PULL_HINT=3"; do
  env $e timeout 600 $TR tools/dist_timeline.py 2>>$o/${tag}_err.log | grep '^{' | tee -a $o/${tag}_timeline_n$N.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['rank'], 'step', round(d['step_ms'], 3), 'whole', round(d['whole_ms'], 3), 'bar', round(d['barrier_ms'], 3), 'pull', round(d['pull_ms'], 3), 'local', round(d['local_ms'], 3), 'remote', round(d['remote_ms'], 3), d['env'])"
done

#!/bin/bash
# round 2, multi-GPU batch ah: timeline of the overlapped step (tools/dist_timeline.py), LDG pull against TMA pull
tag=${1:-r02_ah}; N=${2:-2}; o=gpurun_out; mkdir -p $o
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
for e in "X=1" "GCNB_PULL_TMA=32" "GCNB_PULL_TMA=32 GCNB_PULL_CTAS=2"; do
  env $e timeout 600 $TR tools/dist_timeline.py 2>>$o/${tag}_err.log | grep '^{' | tee -a $o/${tag}_timeline_n$N.jsonl | cut -c1-420
done

#!/bin/bash
# round 2, GPU batch aj: ring depth / occupancy points of the stream kernel on the remote-column block of an 8-way
# partition (latency-bound: every gather is an HBM miss), and on 2- and 4-way partitions
tag=${1:-r02_aj}; o=gpurun_out; mkdir -p $o
for c in 0 1 2 3 4 6 7 9; do
  GCNB_STREAM_CFG=$c python tools/remote_block_probe.py --parts 8 --rank 3 2>>$o/${tag}_err.log | tail -n 1 | python -c "
import sys, json; d = json.loads(sys.stdin.read()); d['stream_cfg'] = $c; print(json.dumps(d))" | tee -a $o/${tag}_remote_cfg.jsonl | python -c "
import sys, json; d = json.loads(sys.stdin.read()); print('cfg', d['stream_cfg'], 'local', round(d['local_ms'], 3), 'remote', round(d['remote_accum_ms'], 3), 'store', round(d['remote_store_ms'], 3))"
done
for c in 0 1 9; do
  GCNB_STREAM_CFG=$c python tools/remote_block_probe.py --parts 2 --rank 1 2>>$o/${tag}_err.log | tail -n 1 | python -c "
import sys, json; d = json.loads(sys.stdin.read()); d['stream_cfg'] = $c; print(json.dumps(d))" | tee -a $o/${tag}_remote_cfg.jsonl | python -c "
import sys, json; d = json.loads(sys.stdin.read()); print('P2 cfg', d['stream_cfg'], 'local', round(d['local_ms'], 3), 'remote', round(d['remote_accum_ms'], 3))"
done

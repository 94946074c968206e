#!/bin/bash
# round 2, GPU batch au: occupancy / ring-depth points of the reworked stream kernel (GCNB_STREAM_CFG) at k = 100 and 128
tag=${1:-r02_au}; o=gpurun_out; mkdir -p $o
B="python bench.py --steps 30 --warmup 5 --comparators none --no-epoch --no-cpu-baseline --no-e2e"
for k in 100 128; do for c in 0 10 11 1 6 3; do
  line=$(GCNB_STREAM_CFG=$c $B --k $k 2>>$o/${tag}_err.log | tail -n 1)
  echo "{\"k\": $k, \"stream_cfg\": $c, \"line\": $line}" >> $o/${tag}_stream_cfg.jsonl
  echo "k=$k cfg=$c $(echo "$line" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d.get("ms_median"), d.get("ms_min"), d["clocks"]["reasons"])')"
done; done

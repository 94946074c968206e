#!/bin/bash
# round 2, GPU batch am: row pitch of B and C against the 128-byte line (k = 48 at 48 / 64 floats, k = 100 at 100 / 112 / 128)
tag=${1:-r02_am}; o=gpurun_out; mkdir -p $o
B="python bench.py --steps 20 --warmup 5 --comparators none --no-epoch --no-cpu-baseline --no-e2e"
for a in "48 48" "48 64" "32 32" "40 40" "40 64" "100 100" "100 112" "100 128" "96 96" "64 64"; do
  set -- $a
  line=$($B --k $1 --pitch $2 2>>$o/${tag}_err.log | tail -n 1)
  echo "{\"k\": $1, \"pitch\": $2, \"line\": $line}" >> $o/${tag}_pitch.jsonl
  echo "k=$1 pitch=$2 $(echo "$line" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d.get("ms_median"), d.get("ms_min"))')"
done

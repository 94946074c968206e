#!/bin/bash
tag=${1:-r01_x}
out=gpurun_out/${tag}_tma_gather_probe.jsonl; : > $out
P="timeout 120 tools/tma_gather_probe"
for cfg in "64 400 8 4 0" "64 400 4 8 0" "64 400 8 2 0" "64 200 8 4 0" "64 512 8 4 0" "64 400 8 4 0 512" "980 400 8 4 4096" "980 400 8 4 0" "980 200 8 4 4096"; do
  $P $cfg >> $out 2>&1
done
cat $out

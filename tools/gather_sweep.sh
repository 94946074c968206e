#!/bin/bash
# sweep of tools/gather_probe; output: gpurun_out/gather_probe.jsonl
out=gpurun_out/gather_probe.jsonl; : > $out
P=tools/gather_probe
for mb in 64 980; do
  for rb in 400 200 512 1024 128; do
    for u in 4 8 16; do
      $P $mb $rb 8 8 $u 0 >> $out
    done
  done
done
# occupancy sensitivity, 400-byte rows out of L2
for cps in 1 2 4; do $P 64 400 8 $cps 8 0 >> $out; done
for cps in 2 4; do $P 64 400 16 $cps 8 0 >> $out; done
# locality window like the products-shaped generator
for u in 4 8 16; do $P 980 400 8 8 $u 4096 >> $out; done
$P 980 200 8 8 8 4096 >> $out
# narrower per-lane loads (8-byte lanes = bf16 VEC4 path)
$P 980 200 8 8 8 4096 8 >> $out
$P 64 200 8 8 8 0 8 >> $out
$P 64 400 8 8 8 0 8 >> $out
cat $out

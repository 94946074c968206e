#!/bin/bash
tag=${1:-r02_j}; o=gpurun_out; mkdir -p $o
out=$o/${tag}_xt_variants.jsonl; : > $out
for v in 0 1 2 3; do GCNB_XT_VARIANT=$v timeout 120 python tools/xt_debug.py >> $out 2>&1; done
GCNB_XW_TALLK_STREAM=0 timeout 120 python tools/xt_debug.py >> $out 2>&1
cut -c1-260 $out

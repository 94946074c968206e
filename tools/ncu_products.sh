#!/bin/bash
# ncu --set full of the products-shaped rows-class kernel for one variant/config
# usage: tools/ncu_products.sh <tag> <kernel-regex> [env assignments...]
tag=$1; shift
kre=$1; shift
env "$@" python bench.py --no-cpu-baseline --no-e2e --steps 2 --warmup 1 > gpurun_out/plain_$tag.log 2>&1 &&
env "$@" ncu --set full --clock-control none --import-source on -k regex:$kre -s 1 -c 1 -o gpurun_out/prof_$tag -f \
    python bench.py --no-cpu-baseline --no-e2e --steps 2 --warmup 1 > gpurun_out/ncu_$tag.log 2>&1

#!/bin/bash
# bench.py over a list of argument strings; prints ms avg / min / roofline frac / GFLOP/s per line
while IFS= read -r args; do
  [ -z "$args" ] && continue
  printf "%-60s " "$args"
  env $ENVV timeout 300 python bench.py --no-cpu-baseline --no-e2e $args 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('avg %.3f min %.3f frac %.4f gflops %.0f' % (d['ms_per_step'], d['ms_min'], d['roofline']['frac'], d['value']))"
done

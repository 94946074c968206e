#!/bin/bash
# R-MAT scale-24 edge-factor-16 feature-width sweep on one B200 (BASELINE config 5, 1-GPU column)
tag=${1:-r01_af}; o=gpurun_out; mkdir -p $o
for k in 32 64 256 512; do
  timeout 400 python bench.py --workload rmat --scale 24 --edge-factor 16 --k $k --no-cpu-baseline --no-e2e --steps 5 --warmup 3 \
     > $o/${tag}_rmat_s24_ef16_k$k.json 2> $o/${tag}_rmat_s24_ef16_k$k.err
  echo -n "k=$k rc=$? "; python - $o/${tag}_rmat_s24_ef16_k$k.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3),"GF",round(d["value"]),"frac",round(d["roofline"]["frac"],3),"nnz",d["config"]["nnz"])
except Exception as e: print("  ERR",e)
PY
  tail -2 $o/${tag}_rmat_s24_ef16_k$k.err | cut -c1-200
done

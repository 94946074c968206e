#!/bin/bash
tag=${1:-r01_t}; o=gpurun_out; mkdir -p $o
run() { name=$1; shift; timeout ${TMO:-600} python bench.py "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo -n "$name rc=$? "; python - $o/${tag}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3))
except Exception as e: print("  ERR",e)
PY
tail -2 $o/${tag}_$name.err; }
Q="--no-cpu-baseline --no-e2e --steps 10 --warmup 3"
for k in 16 32 48 64; do
  run k${k}_rows $Q --k $k
  GCNB_STREAM_MIN_NVEC=3 run k${k}_stream_v4 $Q --k $k
  GCNB_STREAM_MIN_NVEC=3 GCNB_MAX_VEC=2 run k${k}_stream_v2 $Q --k $k
done
GCNB_STREAM_MIN_NVEC=3 run rmat_k32_stream_v4 $Q --workload rmat --scale 22 --edge-factor 16 --k 32
GCNB_STREAM_MIN_NVEC=3 GCNB_MAX_VEC=2 run rmat_k32_stream_v2 $Q --workload rmat --scale 22 --edge-factor 16 --k 32

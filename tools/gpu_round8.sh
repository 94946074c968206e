#!/bin/bash
tag=${1:-r01_u}; o=gpurun_out; mkdir -p $o
run() { name=$1; shift; timeout ${TMO:-600} python bench.py "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo -n "$name rc=$? "; python - $o/${tag}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3))
except Exception as e: print("  ERR",e)
PY
tail -2 $o/${tag}_$name.err; }
Q="--no-cpu-baseline --no-e2e --steps 10 --warmup 3"
export GCNB_STREAM_MIN_NVEC=3
for k in 100 128 256; do GCNB_MAX_VEC=2 run k${k}_stream_v2 $Q --k $k; done
for k in 32 64 100; do GCNB_MAX_VEC=1 run k${k}_stream_v1 $Q --k $k; done
GCNB_MAX_VEC=2 run k24_stream_v2 $Q --k 24
unset GCNB_STREAM_MIN_NVEC
run k24_rows $Q --k 24
GCNB_STREAM_MIN_NVEC=3 GCNB_MAX_VEC=2 run reddit_k128_v2 $Q --workload reddit
GCNB_STREAM_MIN_NVEC=3 GCNB_MAX_VEC=2 run arxiv_k256_v2 $Q --workload arxiv
GCNB_STREAM_MIN_NVEC=3 GCNB_MAX_VEC=2 run arxiv_k128_v2 $Q --workload arxiv --k 128
run arxiv_k128_v4 $Q --workload arxiv --k 128

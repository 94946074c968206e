#!/bin/bash
tag=${1:-r01_v}; o=gpurun_out; mkdir -p $o
run() { name=$1; shift; timeout ${TMO:-600} python bench.py "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo -n "$name rc=$? "; python - $o/${tag}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3))
except Exception as e: print("  ERR",e)
PY
tail -2 $o/${tag}_$name.err; }
timeout 900 python -m pytest tests/test_spmm_gpu.py tests/test_layers_gpu.py tests/test_dist_gpu.py -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -6 $o/${tag}_pytest.log
Q="--no-cpu-baseline --no-e2e --steps 10 --warmup 3"
for k in 8 16 32 48 64; do
  run k${k}_packed $Q --k $k
done
GCNB_PACKED=4 run k16_packed_s4 $Q --k 16
GCNB_PACKED=2 run k16_packed_s2 $Q --k 16
GCNB_PACKED=2 run k32_packed_s2 $Q --k 32
GCNB_PACKED=0 run k8_rows $Q --k 8
run bf16_k64 $Q --k 64 --dtype bf16
run bf16_k128_v4 $Q --k 128 --dtype bf16
GCNB_BF16_V8_MAX=128 run bf16_k128_v8s2 $Q --k 128 --dtype bf16
run rmat_k32 $Q --workload rmat --scale 22 --edge-factor 16 --k 32
run reddit_k48 $Q --workload reddit --k 48
GCNB_PACKED=0 run reddit_k48_rows $Q --workload reddit --k 48
run arxiv_k40 $Q --workload arxiv --k 40
GCNB_PACKED=0 run arxiv_k40_rows $Q --workload arxiv --k 40

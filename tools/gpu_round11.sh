#!/bin/bash
tag=${1:-r01_x}; o=gpurun_out; mkdir -p $o
run() { name=$1; shift; timeout ${TMO:-600} python bench.py "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo -n "$name rc=$? "; python - $o/${tag}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3))
except Exception as e: print("  ERR",e)
PY
tail -2 $o/${tag}_$name.err; }
timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_layers_gpu.py -x -q 2>&1 | tail -2
for w in arxiv products; do
  run epoch_${w}_swap --mode epoch --workload $w --steps 10 --warmup 3
  GCNB_XW_SWAP_DW=0 run epoch_${w}_noswap --mode epoch --workload $w --steps 10 --warmup 3
done
# full ncu captures of the two stream-class kernels on the final code
CMD="python bench.py --no-cpu-baseline --no-e2e --steps 2 --warmup 1"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_stream_kernel -s 1 -c 1 -f -o $o/${tag}_prof_stream_k100 $CMD > $o/${tag}_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_stream_packed -s 1 -c 1 -f -o $o/${tag}_prof_packed_k48 $CMD --k 48 > $o/${tag}_ncu2.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 50 --csv --log-file $o/${tag}_launches_products_k100.csv $CMD > $o/${tag}_ncu3.log 2>&1
ls -la $o | grep ${tag}_prof

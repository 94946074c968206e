#!/bin/bash
tag=${1:-r01_ae}; o=gpurun_out; mkdir -p $o
GCNB_XW_BN256=1 timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_layers_gpu.py -x -q 2>&1 | tail -3
for v in 0 1; do
  GCNB_XW_BN256=$v timeout 600 python bench.py --mode epoch --workload products --steps 10 --warmup 3 --cuda-graph > $o/${tag}_epoch_products_bn256_$v.json 2> $o/${tag}_epoch_products_bn256_$v.err
  echo -n "bn256=$v rc=$? "; python - $o/${tag}_epoch_products_bn256_$v.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3))
except Exception as e: print("  ERR",e)
PY
  tail -2 $o/${tag}_epoch_products_bn256_$v.err | cut -c1-300
done
GCNB_XW_BN256=1 timeout 600 python bench.py --mode epoch --workload arxiv --steps 10 --warmup 3 --cuda-graph | cut -c1-120

#!/bin/bash
tag=${1:-r01_ac}; o=gpurun_out; mkdir -p $o
timeout 600 python -m pytest tests/test_layers_gpu.py -x -q -k "cuda_graph or association" 2>&1 | tail -15
for w in cora arxiv reddit products; do
  timeout 600 python bench.py --mode epoch --workload $w --steps 10 --warmup 3 --cuda-graph > $o/${tag}_epoch_${w}_graph.json 2> $o/${tag}_epoch_${w}_graph.err
  echo -n "$w rc=$? "; python - $o/${tag}_epoch_${w}_graph.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3))
except Exception as e: print("  ERR",e)
PY
  tail -3 $o/${tag}_epoch_${w}_graph.err | cut -c1-300
done

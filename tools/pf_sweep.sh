#!/bin/bash
# software-prefetch modes of the stream kernel: 0 off, 1 L2, 2 L1
tag=${1:-r01_x}
out=gpurun_out/${tag}_pf_sweep.jsonl; : > $out
timeout 600 python -m pytest tests/test_spmm_gpu.py -m gpu -x -q 2>&1 | tail -2
for w in "products f32" "products bf16" "arxiv f32" "reddit f32"; do
  set -- $w
  for pf in 0 1 2; do
    echo "{\"workload\": \"$1\", \"dtype\": \"$2\", \"prefetch\": $pf}" >> $out
    GCNB_PREFETCH=$pf timeout 300 python bench.py --workload $1 --dtype $2 --no-cpu-baseline --no-e2e --steps 10 --warmup 3 >> $out 2>&1
  done
done
grep -o '"workload": "[a-z]*", "dtype": "[a-z0-9]*", "prefetch": [0-9]\|"ms_per_step": [0-9.]*\|"ms_min": [0-9.]*' $out | paste - - -

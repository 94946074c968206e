#!/bin/bash
# stream-kernel configuration sweep on the products-shaped graph; output gpurun_out/<tag>_cfg_sweep.jsonl
tag=${1:-r01_x}
out=gpurun_out/${tag}_cfg_sweep.jsonl; : > $out
for dt in f32 bf16; do
  for cfg in 0 1 2 3 4 5; do
    echo "{\"cfg\": $cfg, \"dtype\": \"$dt\"}" >> $out
    GCNB_STREAM_CFG=$cfg timeout 300 python bench.py --dtype $dt --no-cpu-baseline --no-e2e --steps 10 --warmup 3 >> $out 2>&1
  done
done
for seg in 256 384 1536 3072; do
  echo "{\"seg\": $seg}" >> $out
  GCNB_SEG_NNZ=$seg timeout 300 python bench.py --no-cpu-baseline --no-e2e --steps 10 --warmup 3 >> $out 2>&1
done
python - <<'PY'
import json,sys
tag=None
for l in open(sys.argv[1] if len(sys.argv)>1 else "/dev/stdin"): pass
PY
grep -o '"cfg": [0-9], "dtype": "[a-z0-9]*"\|"seg": [0-9]*\|"ms_per_step": [0-9.]*\|"ms_min": [0-9.]*' $out | paste - - - 

#!/bin/bash
# round 2, GPU batch k: cp.async X loaders in the streaming X·W kernel, tall-K streaming kernel with the accepted
# MN-major layout: parity, per-product times, products epoch
tag=${1:-r02_k}; o=gpurun_out; mkdir -p $o
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_layers_gpu.py -m gpu -x -q > $o/${tag}_pytest.log 2>&1; g=$?; echo "pytest exit $g" >> $o/${tag}_pytest.log; tail -12 $o/${tag}_pytest.log
out=$o/${tag}_xw_shapes.jsonl; : > $out
timeout 300 python tools/xw_time.py >> $out 2>&1
XW_TIME_VARIANTS=1 timeout 300 python tools/xw_time.py >> $out 2>&1
python - <<PY
import json
for l in open("$out"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"{d['ms']:8.3f} ms {d['GBps']:8.1f} GB/s  {d['product'][:48]:48s} {d['env']}")
PY
f=$o/${tag}_epoch_products.json
timeout 600 python bench.py --mode epoch --steps 10 --warmup 3 > $f 2> ${f%.json}.err
echo "epoch products rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"

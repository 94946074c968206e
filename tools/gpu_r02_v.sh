#!/bin/bash
# round 2, GPU batch v: streaming X·W kernel with the A operand (converted X chunks) in tensor memory
tag=${1:-r02_v}; o=gpurun_out; mkdir -p $o
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_layers_gpu.py -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -15 $o/${tag}_pytest.log | cut -c1-200
timeout 300 python tools/xw_time.py > $o/${tag}_xw_shapes.jsonl 2>&1; python - <<PY
import json
for l in open("$o/${tag}_xw_shapes.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"{d['ms']:8.3f} ms {d['GBps']:8.1f} GB/s  {d['product'][:48]:48s}")
PY
XW_TIME_VARIANTS=1 timeout 200 python tools/xw_time.py > $o/${tag}_xw_variants.jsonl 2>&1; python - <<PY
import json
for l in open("$o/${tag}_xw_variants.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"{d['ms']:8.3f} ms  {d['product'][:48]}")
PY
GCNB_XS_TRACE=$o/${tag}_xs_trace.txt XW_TIME_VARIANTS=1 XW_TIME_ONLY="plain    " timeout 200 python tools/xw_time.py 2>&1 | cut -c1-100
python tools/xs_trace_report.py $o/${tag}_xs_trace.txt > $o/${tag}_xs_trace_report.txt 2>&1; rm -f $o/${tag}_xs_trace.txt; grep -v "^    " $o/${tag}_xs_trace_report.txt
f=$o/${tag}_epoch_products.json
timeout 600 python bench.py --mode epoch --workload products --steps 10 --warmup 3 > $f 2> ${f%.json}.err
echo "epoch products rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"

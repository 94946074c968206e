#!/bin/bash
# round 2, GPU batch n: tall-K streaming kernel v2 (bulk row loads + converter re-layout): correctness, per-product
# times, products epoch; default bench line; Reddit ordering sweep at 114.8 M nnz
tag=${1:-r02_n}; o=gpurun_out; mkdir -p $o
timeout 200 python tools/xt_debug.py > $o/${tag}_xt_debug.jsonl 2>&1; echo "xt_debug rc=$?"; cut -c1-160 $o/${tag}_xt_debug.jsonl
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_layers_gpu.py -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -4 $o/${tag}_pytest.log
timeout 300 python tools/xw_time.py > $o/${tag}_xw_shapes.jsonl 2>&1; python - <<PY
import json
for l in open("$o/${tag}_xw_shapes.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"{d['ms']:8.3f} ms {d['GBps']:8.1f} GB/s  {d['product'][:48]:48s}")
PY
for w in products reddit; do
  f=$o/${tag}_epoch_$w.json
  timeout 600 python bench.py --mode epoch --workload $w --steps 10 --warmup 3 > $f 2> ${f%.json}.err
  echo "epoch $w rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
f=$o/${tag}_bench_default.json
s=$(date +%s); timeout 900 python bench.py > $f 2> ${f%.json}.err; echo "default bench rc=$? wall=$(( $(date +%s) - s )) s"
python - <<PY
import json
d=json.load(open("$f"))
print({k:d.get(k) for k in ("value","ms_per_step","ms_median")}, d["roofline"]["frac"], d.get("e2e",{}).get("value"), d.get("epoch"), {k:(v.get("ms") if isinstance(v,dict) else v) for k,v in (d.get("gpu_comparators") or {}).items()}, d.get("cpu_baseline"))
PY
timeout 1500 python tools/reddit_sweep.py > $o/${tag}_reddit_sweep.jsonl 2> $o/${tag}_reddit_sweep.err; echo "reddit sweep rc=$?"; cut -c1-220 $o/${tag}_reddit_sweep.jsonl

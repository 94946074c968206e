#!/bin/bash
# round 2, GPU batch z: full suite on the settled kernels, per-product times, epochs of the four shapes, default bench line
tag=${1:-r02_z}; o=gpurun_out; mkdir -p $o
timeout 1500 python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; g=$?; echo "pytest exit $g" >> $o/${tag}_pytest.log; tail -5 $o/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python tools/xw_time.py > $o/${tag}_xw_shapes.jsonl 2>&1; python - <<PY
import json
for l in open("$o/${tag}_xw_shapes.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"{d['ms']:8.3f} ms {d['GBps']:8.1f} GB/s  {d['product'][:48]:48s}")
PY
for w in products arxiv reddit cora; do
  f=$o/${tag}_epoch_$w.json
  timeout 600 python bench.py --mode epoch --workload $w --steps 10 --warmup 3 > $f 2> ${f%.json}.err
  echo "epoch $w rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
f=$o/${tag}_epoch_cora_graph.json
timeout 600 python bench.py --mode epoch --workload cora --cuda-graph --steps 20 --warmup 5 > $f 2> ${f%.json}.err
echo "epoch cora graph rc=$? $(grep -o '"ms_median": [0-9.]*' $f)"
f=$o/${tag}_bench_default.json
timeout 900 python bench.py > $f 2> ${f%.json}.err; echo "default bench rc=$?"
python - <<PY
import json
d=json.load(open("$f"))
print({k:d[k] for k in ("value","ms_per_step","ms_median")}, d["roofline"]["frac"], d.get("e2e",{}).get("value"), d.get("epoch"), {k:(v.get("ms") if isinstance(v,dict) else v) for k,v in (d.get("gpu_comparators") or {}).items()}, d.get("cpu_baseline"))
PY

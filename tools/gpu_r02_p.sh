#!/bin/bash
# round 2, GPU batch p: tall-K kernel with per-CTA accumulated partials + ones column (bias gradient), accuracy of
# the column sums vs chain length, per-product times, products epoch; Reddit deg order via bench (cached perm)
tag=${1:-r02_p}; o=gpurun_out; mkdir -p $o
for ch in 64 16; do GCNB_XW_MAX_CHAIN=$ch timeout 200 python tools/colsum_probe.py >> $o/${tag}_colsum_probe.jsonl 2>&1; done; cat $o/${tag}_colsum_probe.jsonl | cut -c1-200
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_layers_gpu.py -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -12 $o/${tag}_pytest.log
for ch in 64 16; do
GCNB_XW_MAX_CHAIN=$ch timeout 300 python tools/xw_time.py > $o/${tag}_xw_shapes_chain$ch.jsonl 2>&1; python - <<PY
import json
for l in open("$o/${tag}_xw_shapes_chain$ch.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"chain $ch {d['ms']:8.3f} ms {d['GBps']:8.1f} GB/s  {d['product'][:48]:48s}")
PY
done
f=$o/${tag}_epoch_products.json
timeout 600 python bench.py --mode epoch --workload products --steps 10 --warmup 3 > $f 2> ${f%.json}.err
echo "epoch products rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
CMD="python bench.py --mode epoch --steps 2 --warmup 2"
$CMD > $o/${tag}_plain_epoch.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv \
    --log-file $o/${tag}_launches_epoch_products.csv $CMD > $o/${tag}_ncu_launches.log 2>&1; echo "launch list rc=$?"
python - <<PY
import csv, collections
rows=[r for r in csv.reader(open("$o/${tag}_launches_epoch_products.csv")) if len(r)>10 and r[0].isdigit()]
agg=collections.defaultdict(lambda:[0,0.0])
for r in rows:
    agg[r[4][:60]][0]+=1; agg[r[4][:60]][1]+=float(r[-1].replace(',',''))
for k,(n,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:16]: print(f"{t/1e3:10.1f} us {n:4d}  {k}")
PY
for pf in none profiles/perm/reddit_deg_vomp.npy; do
  f=$o/${tag}_reddit_k128_perm_$(basename $pf .npy).json
  if [ $pf = none ]; then a=""; else a="--perm-file $pf"; fi
  timeout 600 python bench.py --workload reddit --k 128 $a --comparators none --no-epoch --steps 10 --warmup 3 > $f 2> ${f%.json}.err
  echo "reddit k128 $pf rc=$? $(grep -o '"ms_median": [0-9.]*' $f)"
done
f=$o/${tag}_reddit_k128_order_deg.json
timeout 600 python bench.py --workload reddit --k 128 --order deg --comparators none --no-epoch --steps 10 --warmup 3 > $f 2> ${f%.json}.err
echo "reddit k128 order deg (GPU-generated graph) rc=$? $(grep -o '"ms_median": [0-9.]*' $f)"

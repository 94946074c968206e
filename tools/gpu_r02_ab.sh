#!/bin/bash
# round 2, GPU batch ab: faster column-sum fall-back and tall-K reduction; arxiv / products epochs
tag=${1:-r02_ab}; o=gpurun_out; mkdir -p $o
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_layers_gpu.py -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -3 $o/${tag}_pytest.log | cut -c1-200
for w in arxiv products cora; do
  f=$o/${tag}_epoch_$w.json
  timeout 600 python bench.py --mode epoch --workload $w --steps 10 --warmup 3 > $f 2> ${f%.json}.err
  echo "epoch $w rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
CMD="python bench.py --mode epoch --workload arxiv --steps 2 --warmup 2"
$CMD > $o/${tag}_plain_epoch.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 300 --csv \
    --log-file $o/${tag}_launches_epoch_arxiv.csv $CMD > $o/${tag}_ncu_launches.log 2>&1; echo "launch list rc=$?"
python - <<PY
import csv, collections
rows=[r for r in csv.reader(open("$o/${tag}_launches_epoch_arxiv.csv")) if len(r)>10 and r[0].isdigit()]
agg=collections.defaultdict(lambda:[0,0.0])
for r in rows[-120:]:
    agg[r[4][:70]][0]+=1; agg[r[4][:70]][1]+=float(r[-1].replace(',',''))
for k,(n,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:20]: print(f"{t/1e3:10.1f} us {n:4d}  {k}")
PY
CMD="python bench.py --mode epoch --steps 2 --warmup 2"
$CMD > $o/${tag}_plain_epoch_products.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 400 --csv \
    --log-file $o/${tag}_launches_epoch_products.csv $CMD > $o/${tag}_ncu_launches2.log 2>&1; echo "launch list rc=$?"
python - <<PY
import csv, collections
rows=[r for r in csv.reader(open("$o/${tag}_launches_epoch_products.csv")) if len(r)>10 and r[0].isdigit()]
agg=collections.defaultdict(lambda:[0,0.0])
for r in rows[-64:]:
    agg[r[4][:70]][0]+=1; agg[r[4][:70]][1]+=float(r[-1].replace(',',''))
for k,(n,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:20]: print(f"{t/1e3:10.1f} us {n:4d}  {k}")
PY

#!/bin/bash
# round 2, GPU batch aa: why did the arxiv-shaped epoch get slower (1.29 -> 1.81 ms)?  per-product times at n = 169 343
# and the launch list of that epoch
tag=${1:-r02_aa}; o=gpurun_out; mkdir -p $o
for kb in 128 64; do
GCNB_XS_WRES_KB=$kb timeout 300 python tools/xw_time.py 169343 > $o/${tag}_xw_shapes_arxiv_wres$kb.jsonl 2>&1; python - <<PY
import json
for l in open("$o/${tag}_xw_shapes_arxiv_wres$kb.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"wres_kb $kb {d['ms']:8.4f} ms {d['GBps']:8.1f} GB/s  {d['product'][:48]:48s}")
PY
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

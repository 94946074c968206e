#!/bin/bash
# round 2, GPU batch o: launch list of the products-shaped epoch and a source-level profile of the tall-K
# streaming kernel (v2, bulk row loads) — what bounds it now that the TMA row rate is out of the picture?
tag=${1:-r02_o}; o=gpurun_out; mkdir -p $o
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
$CMD > /dev/null 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:xt_kernel -s 2 -c 2 -f -o /tmp/prof_xt $CMD > $o/${tag}_ncu_xt.log 2>&1
echo "ncu xt rc=$?"
ncu -i /tmp/prof_xt.ncu-rep --page source --csv > /tmp/xt_source.csv 2>/dev/null
python tools/ncu_source_top.py /tmp/xt_source.csv > $o/${tag}_ncu_xt_source_top.txt 2>&1
ncu -i /tmp/prof_xt.ncu-rep --page details --csv 2>/dev/null | python -c "
import csv,sys
for r in csv.reader(sys.stdin):
    if len(r)>14 and any(k in r[12] for k in ('Duration','Throughput','Stall','Eligible','Issued','Active Warps','Registers','Hit Rate','Pipe','Bank')): print(r[4][:40],'|',r[11][:30],'|',r[12][:50],'|',r[13],'|',r[14])
" > $o/${tag}_ncu_xt_details.txt 2>&1
python profiles/ncu_summary.py /tmp/prof_xt.ncu-rep dram__bytes sm__pipe_tensor l1tex__data_bank smsp__warp_issue_stalled l1tex__data_pipe > $o/${tag}_ncu_xt_kernel.txt 2>&1
rm -f /tmp/prof_xt.ncu-rep; du -sh $o; head -70 $o/${tag}_ncu_xt_source_top.txt; grep -i "duration\|dram__bytes" $o/${tag}_ncu_xt_kernel.txt | head

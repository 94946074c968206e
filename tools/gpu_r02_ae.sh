#!/bin/bash
# round 2, GPU batch ae: source-level profile of the packed stream kernel at k = 48 (half of the products epoch) and
# of the stream kernel at k = 100 (the headline): where do the instructions go?
tag=${1:-r02_ae}; o=gpurun_out; mkdir -p $o
for k in 48 100; do
  CMD="python bench.py --k $k --steps 2 --warmup 1 --comparators none --no-epoch"
  $CMD > $o/${tag}_plain_k$k.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_stream -s 1 -c 1 -f -o /tmp/prof_k$k $CMD > $o/${tag}_ncu_k$k.log 2>&1
  echo "ncu k=$k rc=$?"
  ncu -i /tmp/prof_k$k.ncu-rep --page source --csv > /tmp/src_k$k.csv 2>/dev/null
  python tools/ncu_source_top.py /tmp/src_k$k.csv 70 > $o/${tag}_ncu_spmm_k${k}_source_top.txt 2>&1
  python profiles/ncu_summary.py /tmp/prof_k$k.ncu-rep dram__bytes smsp__inst_executed l1tex__data_pipe smsp__warp_issue_stalled smsp__issue_active sm__inst_executed_pipe > $o/${tag}_ncu_spmm_k${k}.txt 2>&1
  rm -f /tmp/prof_k$k.ncu-rep
done
du -sh $o; head -90 $o/${tag}_ncu_spmm_k48_source_top.txt | cut -c1-170

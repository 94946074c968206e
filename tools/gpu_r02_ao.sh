#!/bin/bash
# round 2, GPU batch ao: source-level ncu of the reworked packed (k = 48) and stream (k = 100) kernels
tag=${1:-r02_ao}; o=gpurun_out; mkdir -p $o
for k in 48 100; do
  CMD="python bench.py --k $k --steps 2 --warmup 1 --comparators none --no-epoch --no-cpu-baseline --no-e2e"
  $CMD > $o/${tag}_plain_k$k.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm_stream -s 1 -c 1 -f -o /tmp/prof_k$k $CMD > $o/${tag}_ncu_k$k.log 2>&1
  echo "ncu k=$k rc=$?"
  ncu -i /tmp/prof_k$k.ncu-rep --page source --csv > /tmp/src_k$k.csv 2>/dev/null
  python tools/ncu_source_top.py /tmp/src_k$k.csv 60 > $o/${tag}_ncu_spmm_k${k}_source_top.txt 2>&1
  python profiles/ncu_summary.py /tmp/prof_k$k.ncu-rep gpu__time dram__bytes_read.sum dram__bytes_write.sum sm__inst_executed.sum l1tex__data_pipe_lsu_wavefronts smsp__issue_active lts__t_sector_hit_rate lts__t_sectors_srcunit_tex.sum l1tex__m_xbar2l1tex_read_bytes launch__registers sm__warps_active > $o/${tag}_ncu_spmm_k${k}.txt 2>&1
  rm -f /tmp/prof_k$k.ncu-rep
done
head -64 $o/${tag}_ncu_spmm_k48_source_top.txt | cut -c1-150

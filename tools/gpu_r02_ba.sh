#!/bin/bash
# round 2, GPU batch ba: ncu --set full of the shipped stream kernel (k = 100) and launch list of the default bench
tag=${1:-r02_ba}; o=gpurun_out; mkdir -p $o
CMD="python bench.py --k 100 --steps 2 --warmup 1 --comparators none --no-epoch --no-cpu-baseline --no-e2e"
$CMD > $o/${tag}_plain.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:spmm_stream -s 1 -c 1 -f -o /tmp/prof_k100 $CMD > $o/${tag}_ncu.log 2>&1
echo "ncu rc=$?"
python profiles/ncu_summary.py /tmp/prof_k100.ncu-rep gpu__time dram__bytes_read.sum dram__bytes_write.sum sm__inst_executed.sum l1tex__data_pipe_lsu_wavefronts smsp__issue_active lts__t_sector_hit_rate l1tex__m_xbar2l1tex_read_bytes launch__registers launch__occupancy sm__warps_active smsp__average_warps_issue_stalled_long dram__throughput > $o/${tag}_ncu_spmm_k100.txt 2>&1
grep -E "gpu__time_duration.sum|dram__bytes_read.sum \[|dram__bytes_write.sum \[|lsu_wavefronts.avg.pct|registers_per_thread \[|warps_active.avg.pct" $o/${tag}_ncu_spmm_k100.txt
CMD2="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --comparators none"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $o/${tag}_launches_bench_default.csv $CMD2 > $o/${tag}_ncu_launches.log 2>&1; echo "launch list rc=$?"

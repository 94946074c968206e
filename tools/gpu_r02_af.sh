#!/bin/bash
# round 2, GPU batch af: the blocks of one rank of the 8-way (and 2-way) row partition timed alone on one GPU
tag=${1:-r02_af}; o=gpurun_out; mkdir -p $o
for args in "--parts 8 --rank 3" "--parts 8 --rank 3 --seg-nnz 32" "--parts 8 --rank 3 --seg-nnz 64" "--parts 8 --rank 3 --seg-nnz 512" "--parts 8 --rank 0" "--parts 2 --rank 1" "--parts 8 --rank 3 --k 48"; do
  python tools/remote_block_probe.py $args 2>>$o/${tag}_err.log | tail -n 1 | tee -a $o/${tag}_remote_block.jsonl
done
CMD="python tools/remote_block_probe.py --parts 8 --rank 3 --iters 2"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmm_stream -s 12 -c 1 -f -o /tmp/prof_rb $CMD > $o/${tag}_ncu.log 2>&1
echo "ncu rc=$?"
ncu -i /tmp/prof_rb.ncu-rep --page source --csv > /tmp/src_rb.csv 2>/dev/null
python tools/ncu_source_top.py /tmp/src_rb.csv 50 > $o/${tag}_ncu_remote_source_top.txt 2>&1
python profiles/ncu_summary.py /tmp/prof_rb.ncu-rep gpu__time dram__bytes_read.sum dram__bytes_write.sum smsp__inst_executed.sum l1tex__data_pipe_lsu_wavefronts.avg.pct smsp__issue_active sm__warps_active launch__grid launch__registers > $o/${tag}_ncu_remote.txt 2>&1
head -40 $o/${tag}_ncu_remote_source_top.txt | cut -c1-160; cat $o/${tag}_ncu_remote.txt | head -30

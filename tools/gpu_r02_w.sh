#!/bin/bash
# round 2, GPU batch w: finer timeline of the MMA warp of the streaming X·W kernel
tag=${1:-r02_w}; o=gpurun_out; mkdir -p $o
GCNB_XS_TRACE=$o/${tag}_xs_trace.txt XW_TIME_VARIANTS=1 XW_TIME_ONLY="plain    " timeout 200 python tools/xw_time.py 2>&1 | cut -c1-100
python tools/xs_trace_report.py $o/${tag}_xs_trace.txt > $o/${tag}_xs_trace_report.txt 2>&1; rm -f $o/${tag}_xs_trace.txt; cat $o/${tag}_xs_trace_report.txt

#!/bin/bash
# round 2, GPU batch t: clock64() timeline of the streaming X·W kernel's roles (CTA 0) on the layer-1 product
tag=${1:-r02_t}; o=gpurun_out; mkdir -p $o
for dbg in 0 3; do
  GCNB_XS_TRACE=$o/${tag}_xs_trace_dbg$dbg.txt GCNB_XS_DEBUG=$dbg XW_TIME_VARIANTS=1 XW_TIME_ONLY="plain    " timeout 200 python tools/xw_time.py 2>&1 | cut -c1-120
done
ls -la $o/${tag}_xs_trace_*

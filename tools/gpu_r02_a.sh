#!/bin/bash
# round 2, GPU batch a: regression suite after the L2-policy change, tensor-map gather4 probe, L2 policy A/B
tag=${1:-r02_a}; o=gpurun_out; mkdir -p $o
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
timeout 900 python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -3 $o/${tag}_pytest.log
P="timeout 120 tools/tma_gather4_probe"
out=$o/${tag}_tma_gather4_probe.jsonl; : > $out
#          table_mb row_bytes consumer_warps ctas_per_sm stages window pitch
for cfg in "64 400 8 1 0" "64 400 16 1 0" "64 400 4 1 0" "64 400 8 2 0" "64 400 8 1 4" "64 400 8 1 8" \
           "64 200 8 1 0" "64 512 8 1 0" "64 256 8 1 0" "64 1024 8 1 0" \
           "980 400 8 1 0 4096" "980 400 16 1 0 4096" "980 400 8 2 0 4096" "980 400 8 1 0 0" "980 512 8 1 0 4096" "980 200 8 1 0 4096"; do
  $P $cfg >> $out 2>&1
done
cat $out | cut -c1-330
for pol in 0 1 2 3; do
  GCNB_L2POL=$pol timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 20 --warmup 5 > $o/${tag}_bench_l2pol$pol.json 2> $o/${tag}_bench_l2pol$pol.err
  echo "l2pol=$pol rc=$? $(grep -o '"ms_per_step": [0-9.]*\|"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $o/${tag}_bench_l2pol$pol.json | paste - - -)"
done
for seg in 384 1536; do
  GCNB_SEG_NNZ=$seg timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 20 --warmup 5 > $o/${tag}_bench_seg$seg.json 2> $o/${tag}_bench_seg$seg.err
  echo "seg=$seg rc=$? $(grep -o '"ms_per_step": [0-9.]*\|"ms_median": [0-9.]*' $o/${tag}_bench_seg$seg.json | paste - -)"
done

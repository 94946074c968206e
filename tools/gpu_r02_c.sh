#!/bin/bash
# round 2, GPU batch c: stream-kernel ring/occupancy points with U = 3, ncu --set full of the products stream
# kernel and of the Reddit-shaped stream / panel kernels (L1 hit rates, L2->SM bytes)
tag=${1:-r02_c}; o=gpurun_out; mkdir -p $o
Q="--no-cpu-baseline --no-e2e --no-epoch --comparators none"
for cfg in 0 6 7 8 9 4; do
  f=$o/${tag}_products_cfg$cfg.json
  GCNB_STREAM_CFG=$cfg timeout 600 python bench.py $Q --steps 20 --warmup 5 > $f 2> ${f%.json}.err
  echo "products cfg=$cfg rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
for cfg in 0 6 8; do
  f=$o/${tag}_reddit_deg_cfg$cfg.json
  GCNB_PANEL=0 GCNB_STREAM_CFG=$cfg timeout 600 python bench.py --workload reddit --k 128 --order deg $Q --steps 10 --warmup 3 > $f 2> ${f%.json}.err
  echo "reddit deg cfg=$cfg rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
N="--steps 2 --warmup 1"
prof() {  # tag kernel-regex env... -- bench args
  t=$1; k=$2; shift 2
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" python bench.py $Q $N "$@" > $o/${tag}_plain_$t.log 2>&1 &&
  env "${envs[@]}" timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -f -o $o/${tag}_prof_$t \
      python bench.py $Q $N "$@" > $o/${tag}_ncu_$t.log 2>&1
  echo "ncu $t rc=$?"
}
prof products_stream spmm_stream_kernel GCNB_PANEL=0 --
prof reddit_deg_stream spmm_stream_kernel GCNB_PANEL=0 -- --workload reddit --k 128 --order deg
prof reddit_deg_panel spmm_panel_kernel GCNB_PANEL=1 GCNB_PANEL_TILE=2048 -- --workload reddit --k 128 --order deg
prof reddit_deg_k602 spmm_stream_kernel GCNB_PANEL=0 -- --workload reddit --k 602 --order deg
ls -la $o/${tag}_prof_* 2>/dev/null

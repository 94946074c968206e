#!/bin/bash
# round 2, GPU batch e: full GPU suite, streaming X·W kernel v2 + fused head in the products epoch, Reddit k=602
# column slices, R-MAT scale-24 edge-factor 32 / 64 width sweeps
tag=${1:-r02_e}; o=gpurun_out; mkdir -p $o
timeout 1500 python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; g=$?; echo "pytest exit $g" >> $o/${tag}_pytest.log; tail -12 $o/${tag}_pytest.log
for st in 1 0; do
  f=$o/${tag}_epoch_products_stream$st.json
  GCNB_XW_STREAM=$st timeout 600 python bench.py --mode epoch --steps 10 --warmup 3 > $f 2> ${f%.json}.err
  echo "epoch products xw_stream=$st rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
CMD="python bench.py --mode epoch --steps 2 --warmup 2"
$CMD > $o/${tag}_plain_epoch.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv \
    --log-file $o/${tag}_launches_epoch_products.csv $CMD > $o/${tag}_ncu_launches.log 2>&1; echo "launch list rc=$?"
Q="--no-cpu-baseline --no-e2e --no-epoch --comparators none --steps 10 --warmup 3"
for nv in 0 4; do
  f=$o/${tag}_reddit_k602_deg_widenv$nv.json
  GCNB_PANEL=0 GCNB_WIDE_NV=$nv timeout 900 python bench.py --workload reddit --k 602 --order deg $Q > $f 2> ${f%.json}.err
  echo "reddit k602 deg wide_nv=$nv rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
for pol in 1 3; do
  f=$o/${tag}_reddit_k128_deg_l2pol$pol.json
  GCNB_PANEL=0 GCNB_L2POL=$pol timeout 900 python bench.py --workload reddit --k 128 --order deg $Q > $f 2> ${f%.json}.err
  echo "reddit k128 deg l2pol=$pol rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
for ef in 32 64; do
  f=$o/${tag}_rmat_s24_ef${ef}_ksweep.jsonl
  timeout 1500 python bench.py --workload rmat --scale 24 --edge-factor $ef --k-sweep 32,128,512 --steps 5 --warmup 3 > $f 2> ${f%.jsonl}.err
  echo "rmat s24 ef$ef rc=$?"; grep -o '"k": [0-9]*\|"nnz": [0-9]*\|"ms_median": [0-9.]*\|"frac": [0-9.]*\|"checksum_rel_err": [0-9.e-]*' $f | paste - - - - -; tail -2 ${f%.jsonl}.err
done
rm -f $o/*.ncu-rep; du -sh $o

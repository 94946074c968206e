#!/bin/bash
# round 2, GPU batch f: streaming X·W kernel v3 (resident W), faster loss kernels, products epoch + launch list,
# ncu of the streaming kernel, R-MAT scale-24 edge-factor 64
tag=${1:-r02_f}; o=gpurun_out; mkdir -p $o
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_layers_gpu.py -m gpu -x -q > $o/${tag}_pytest.log 2>&1; g=$?; echo "pytest exit $g" >> $o/${tag}_pytest.log; tail -8 $o/${tag}_pytest.log
for wr in 1 0; do
  f=$o/${tag}_epoch_products_wres$wr.json
  GCNB_XW_WRES=$wr timeout 600 python bench.py --mode epoch --steps 10 --warmup 3 > $f 2> ${f%.json}.err
  echo "epoch products wres=$wr rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
for w in arxiv reddit cora; do
  f=$o/${tag}_epoch_$w.json
  timeout 600 python bench.py --mode epoch --workload $w --steps 10 --warmup 3 > $f 2> ${f%.json}.err
  echo "epoch $w rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
CMD="python bench.py --mode epoch --steps 2 --warmup 2"
$CMD > $o/${tag}_plain_epoch.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv \
    --log-file $o/${tag}_launches_epoch_products.csv $CMD > $o/${tag}_ncu_launches.log 2>&1; echo "launch list rc=$?"
$CMD > /dev/null 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:xs_kernel -s 4 -c 3 -f -o /tmp/prof_xs $CMD > $o/${tag}_ncu_xs.log 2>&1
echo "ncu xs rc=$?"
python profiles/ncu_summary.py /tmp/prof_xs.ncu-rep sm__pipe_tensor sm__inst_executed_pipe_uniform smsp__warp_issue_stalled > $o/${tag}_ncu_xs_kernel_epoch_products.txt 2>&1
rm -f /tmp/prof_xs.ncu-rep
f=$o/${tag}_rmat_s24_ef64_ksweep.jsonl
timeout 1800 python bench.py --workload rmat --scale 24 --edge-factor 64 --k-sweep 32,128,512 --steps 5 --warmup 3 > $f 2> ${f%.jsonl}.err
echo "rmat s24 ef64 rc=$?"; grep -o '"k": [0-9]*\|"nnz": [0-9]*\|"ms_median": [0-9.]*\|"frac": [0-9.]*\|"checksum_rel_err": [0-9.e-]*' $f | paste - - - - -; tail -2 ${f%.jsonl}.err
rm -f $o/*.ncu-rep; du -sh $o

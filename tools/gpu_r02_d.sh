#!/bin/bash
# round 2, GPU batch d: streaming X·W kernel parity, products epoch A/B, ncu summaries (made on the box: the
# .ncu-rep files are too large to bring back)
tag=${1:-r02_d}; o=gpurun_out; mkdir -p $o
timeout 420 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q > $o/${tag}_pytest_gemm.log 2>&1; g=$?; echo "pytest gemm exit $g" >> $o/${tag}_pytest_gemm.log; tail -15 $o/${tag}_pytest_gemm.log
if [ $g -ne 0 ]; then export GCNB_XW_STREAM=0; echo "streaming kernel disabled for the rest of this batch"; fi
timeout 600 python -m pytest tests/test_layers_gpu.py tests/test_dist_gpu.py -m gpu -x -q > $o/${tag}_pytest_layers.log 2>&1; echo "pytest layers/dist exit $?" >> $o/${tag}_pytest_layers.log; tail -4 $o/${tag}_pytest_layers.log
for st in 1 0; do
  f=$o/${tag}_epoch_products_stream$st.json
  GCNB_XW_STREAM=${GCNB_XW_STREAM:-$st} timeout 600 python bench.py --mode epoch --steps 10 --warmup 3 > $f 2> ${f%.json}.err
  echo "epoch products xw_stream=$st rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
CMD="python bench.py --mode epoch --steps 2 --warmup 2"
$CMD > $o/${tag}_plain_epoch.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv \
    --log-file $o/${tag}_launches_epoch_products.csv $CMD > $o/${tag}_ncu_launches.log 2>&1; echo "launch list rc=$?"
Q="--no-cpu-baseline --no-e2e --no-epoch --comparators none --steps 2 --warmup 1"
prof() {  # tag kernel-regex env... -- bench args
  t=$1; k=$2; shift 2
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" python bench.py $Q "$@" > $o/${tag}_plain_$t.log 2>&1 &&
  env "${envs[@]}" timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -f -o /tmp/prof_$t \
      python bench.py $Q "$@" > $o/${tag}_ncu_$t.log 2>&1
  echo "ncu $t rc=$?"
  python profiles/ncu_summary.py /tmp/prof_$t.ncu-rep l1tex__m_xbar2l1tex l1tex__t_sector_hit l1tex__data_pipe smsp__pcsamp sm__inst_executed_pipe_lsu \
      l1tex__lsu_writeback l1tex__m_l1tex2xbar lts__t_sectors_srcunit_tex_op_read l1tex__t_sectors_pipe_lsu_mem_global_op_ld_lookup > $o/${tag}_ncu_$t.txt 2>&1
  ncu -i /tmp/prof_$t.ncu-rep --page details --csv 2>/dev/null | grep -i -E "stall|hit rate|throughput|Issued|No Eligible|L2 Cache|Mem Busy|Max Bandwidth|Registers|Occupancy|Duration" | cut -c1-300 > $o/${tag}_ncu_${t}_details.txt
  rm -f /tmp/prof_$t.ncu-rep
}
prof products_stream spmm_stream_kernel GCNB_PANEL=0 --
prof reddit_deg_stream spmm_stream_kernel GCNB_PANEL=0 -- --workload reddit --k 128 --order deg
prof reddit_deg_panel spmm_panel_kernel GCNB_PANEL=1 GCNB_PANEL_TILE=2048 -- --workload reddit --k 128 --order deg
prof reddit_deg_k602 spmm_stream_kernel GCNB_PANEL=0 -- --workload reddit --k 602 --order deg
du -sh $o

#!/bin/bash
# round 2, GPU batch an: stream + packed kernels with the cp.async slab ring, unpredicated gathers and whole-quarter
# turns (GCNB_PACKED_MEDIUM=0 switches the whole-quarter / cheap row-end paths off): parity, then the width sweep
tag=${1:-r02_an}; o=gpurun_out; mkdir -p $o
python -m pytest tests/test_spmm_gpu.py tests/test_layers_gpu.py tests/test_dist_gpu.py -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest rc=$? $(tail -n 1 $o/${tag}_pytest.log)"
B="python bench.py --steps 20 --warmup 5 --comparators none --no-epoch --no-cpu-baseline --no-e2e"
for k in 48 100 128 256; do for m in 0 1; do
  line=$(GCNB_PACKED_MEDIUM=$m $B --k $k 2>>$o/${tag}_err.log | tail -n 1)
  echo "{\"k\": $k, \"medium\": $m, \"line\": $line}" >> $o/${tag}_sweep.jsonl
  echo "k=$k medium=$m $(echo "$line" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d.get("ms_median"), d.get("ms_min"))')"
done; done
for w in "reddit 128" "reddit 602" "arxiv 128"; do set -- $w
  line=$($B --workload $1 --k $2 2>>$o/${tag}_err.log | tail -n 1)
  echo "{\"workload\": \"$1\", \"k\": $2, \"line\": $line}" >> $o/${tag}_sweep.jsonl
  echo "$1 k=$2 $(echo "$line" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d.get("ms_median"), d.get("ms_min"))')"
done
python bench.py --mode epoch --steps 10 --warmup 3 --no-cpu-baseline > $o/${tag}_epoch_products.json 2>>$o/${tag}_err.log
echo "epoch $(grep -o '"ms_per_step": [0-9.]*' $o/${tag}_epoch_products.json)"

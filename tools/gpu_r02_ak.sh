#!/bin/bash
# round 2, GPU batch ak: packed stream kernel with a cheaper path for ring turns that contain a row end
# (GCNB_PACKED_MEDIUM), parity first, then k = 16..64 on the products shape and the products epoch
tag=${1:-r02_ak}; o=gpurun_out; mkdir -p $o
python -m pytest tests/test_spmm_gpu.py tests/test_layers_gpu.py tests/test_dist_gpu.py -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest rc=$? $(tail -n 1 $o/${tag}_pytest.log)"
B="python bench.py --steps 20 --warmup 5 --comparators none --no-epoch --no-cpu-baseline --no-e2e"
for k in 16 32 48 64; do for m in 0 1; do
  line=$(GCNB_PACKED_MEDIUM=$m $B --k $k 2>>$o/${tag}_err.log | tail -n 1)
  echo "{\"k\": $k, \"medium\": $m, \"line\": $line}" >> $o/${tag}_packed_medium.jsonl
  echo "k=$k medium=$m $(echo "$line" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d.get("ms_median"), d.get("ms_min"))')"
done; done
for m in 0 1; do
  GCNB_PACKED_MEDIUM=$m python bench.py --mode epoch --steps 10 --warmup 3 --no-cpu-baseline > $o/${tag}_epoch_products_medium$m.json 2>>$o/${tag}_err.log
  echo "epoch medium=$m $(grep -o '"ms_per_step": [0-9.]*' $o/${tag}_epoch_products_medium$m.json)"
done

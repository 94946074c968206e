#!/bin/bash
# round 2, GPU batch b: panel-class parity, Reddit-shaped (114.8 M nnz) panel vs stream, comparators
tag=${1:-r02_b}; o=gpurun_out; mkdir -p $o
timeout 1200 python -m pytest tests/test_spmm_gpu.py tests/test_layers_gpu.py -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -4 $o/${tag}_pytest.log
Q="--no-cpu-baseline --no-e2e --no-epoch --comparators none --steps 10 --warmup 3"
run() {  # name, env..., -- args
  name=$1; shift
  env "$@" > /dev/null 2>&1
}
for ord in none deg; do
  for panel in 0 1; do
    f=$o/${tag}_reddit_k128_${ord}_panel$panel.json
    GCNB_PANEL=$panel timeout 900 python bench.py --workload reddit --k 128 --order $ord $Q > $f 2> ${f%.json}.err
    echo "reddit k128 order=$ord panel=$panel rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*\|"nnz": [0-9]*' $f | paste - - -)"
  done
done
for tile in 256 1024 2048; do
  f=$o/${tag}_reddit_k128_deg_panel1_tile$tile.json
  GCNB_PANEL=1 GCNB_PANEL_TILE=$tile timeout 900 python bench.py --workload reddit --k 128 --order deg $Q > $f 2> ${f%.json}.err
  echo "reddit k128 deg tile=$tile rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
for panel in 0 1; do
  f=$o/${tag}_reddit_k602_deg_panel$panel.json
  GCNB_PANEL=$panel timeout 900 python bench.py --workload reddit --k 602 --order deg $Q > $f 2> ${f%.json}.err
  echo "reddit k602 deg panel=$panel rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
done
f=$o/${tag}_products_comparators.json
timeout 1200 python bench.py --comparators all --no-epoch --no-cpu-baseline --no-e2e > $f 2> ${f%.json}.err
echo "products comparators rc=$?"; python - <<PY
import json
d=json.load(open("$f")); print(d["ms_median"]); print(json.dumps(d.get("gpu_comparators"), indent=1)[:3000])
PY

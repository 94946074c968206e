#!/bin/bash
tag=${1:-r01_s}; o=gpurun_out; mkdir -p $o
run() { name=$1; shift; timeout ${TMO:-600} python bench.py "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo "$name rc=$?"; python - $o/${tag}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("  ",d["metric"],"avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3), d["config"].get("association"), d["clocks"])
except Exception as e: print("  ERR",e)
PY
tail -2 $o/${tag}_$name.err; }
timeout 900 python -m pytest tests/test_layers_gpu.py -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -5 $o/${tag}_pytest.log
for w in cora arxiv reddit products; do for a in auto reference; do run epoch_${w}_$a --mode epoch --workload $w --assoc $a --steps 10 --warmup 3; done; done
run products_k48 --k 48 --no-cpu-baseline --no-e2e

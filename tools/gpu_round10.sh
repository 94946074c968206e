#!/bin/bash
tag=${1:-r01_w}; o=gpurun_out; mkdir -p $o
run() { name=$1; shift; timeout ${TMO:-600} python bench.py "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo -n "$name rc=$? "; python - $o/${tag}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3))
except Exception as e: print("  ERR",e)
PY
tail -2 $o/${tag}_$name.err; }
timeout 1500 python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -4 $o/${tag}_pytest.log
for w in arxiv reddit products; do run epoch_${w}_auto --mode epoch --workload $w --steps 10 --warmup 3; done
CMD="python bench.py --mode epoch --workload products --steps 2 --warmup 1"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file $o/${tag}_launches_epoch_products.csv $CMD > $o/${tag}_ncu_launches.log 2>&1
tail -2 $o/${tag}_ncu_launches.log

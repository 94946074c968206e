#!/bin/bash
# Reddit-shaped reorder sweep with permutations computed once on the host (profiles/perm/, host seconds in
# reddit_host_seconds.json); the graph is generated on the CPU so that it is the one the files belong to
tag=${1:-r01_aa}; o=gpurun_out; mkdir -p $o
Q="--no-cpu-baseline --no-e2e --steps 20 --warmup 5 --workload reddit"
for ord in identity deg rcm rabbit gorder; do
  timeout 600 python bench.py $Q --perm-file profiles/perm/reddit_${ord}_vomp.npy > $o/${tag}_reddit_$ord.json 2> $o/${tag}_reddit_$ord.err
  echo -n "$ord rc=$? "; python - $o/${tag}_reddit_$ord.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3), d["config"]["order"])
except Exception as e: print("  ERR",e)
PY
  tail -2 $o/${tag}_reddit_$ord.err
done
timeout 600 python bench.py $Q --k 602 > $o/${tag}_reddit_k602.json 2> $o/${tag}_reddit_k602.err; cut -c1-200 $o/${tag}_reddit_k602.json
timeout 600 python bench.py --no-cpu-baseline --no-e2e --order rabbit > $o/${tag}_products_rabbit.json 2> $o/${tag}_products_rabbit.err; cut -c1-700 $o/${tag}_products_rabbit.json | cut -c150-700

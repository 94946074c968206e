#!/bin/bash
tag=${1:-r01_q}; o=gpurun_out; mkdir -p $o
runm() { name=$1; shift; NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
  --master-port 29511 bench.py --gpus 2 "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo "$name rc=$?";
  python - $o/${tag}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3),"min",round(d["ms_min"],3), d.get("ms_steps_rank0"))
except Exception as e: print("  ERR", e)
PY
}
runm pull_s10 --steps 10 --warmup 3
runm pull_s20 --steps 20 --warmup 5

#!/bin/bash
tag=${1:-r01_y}; o=gpurun_out; mkdir -p $o
runm() { n=$1; name=$2; shift 2; NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
  --master-port $((29520 + n)) bench.py --gpus $n --steps 20 --warmup 5 "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo "$name rc=$?";
  python - $o/${tag}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3),"min",round(d["ms_min"],3), d.get("phase_ms_alone"), d.get("ms_steps_rank0"))
except Exception as e: print("  ERR", e)
PY
  tail -2 $o/${tag}_$name.err | cut -c1-300
}
runm 8 spmm_n8 
runm 8 epoch_n8 --mode epoch
runm 4 epoch_n4 --mode epoch
runm 2 epoch_n2 --mode epoch
runm 4 spmm_n4

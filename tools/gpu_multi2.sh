#!/bin/bash
# 2-GPU experiments on the exchange: pull-kernel footprint, exchange mode
tag=${1:-r01_m}; o=gpurun_out; mkdir -p $o
timeout 600 python -m pytest tests/test_dist_gpu.py tests/test_spmm_gpu.py -x -q 2>&1 | tail -3
runm() { name=$1; shift; NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
  --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo "$name rc=$?"; 
  python - $o/${tag}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3),"min",round(d["ms_min"],3), d.get("phase_ms_alone"))
except Exception as e: print("  ERR", e)
PY
}
for c in 1 2 4; do GCNB_PULL_CTAS=$c runm pull_ctas$c --exchange pull; done
runm allgather --exchange allgather
runm epoch_pull --exchange pull --mode epoch
runm epoch_allgather --exchange allgather --mode epoch
runm arxiv_pull --exchange pull --workload arxiv
runm reddit_pull --exchange pull --workload reddit

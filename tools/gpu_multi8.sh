#!/bin/bash
# 8-GPU box: strong scaling of the row-partitioned SpMM / GCN epoch at 4 and 8 ranks
tag=${1:-r01_o}; o=gpurun_out; mkdir -p $o
nvidia-smi topo -m > $o/${tag}_topo.txt 2>&1
runm() { n=$1; name=$2; shift 2; NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
  --master-port $((29520 + n)) bench.py --gpus $n --steps 10 --warmup 3 "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo "$name rc=$?";
  python - $o/${tag}_$name.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("  avg",round(d["ms_per_step"],3),"med",round(d["ms_median"],3),"min",round(d["ms_min"],3), d.get("phase_ms_alone"), d["config"].get("exchange_bytes_received_per_rank_max"), d["config"].get("local_nnz_frac"))
except Exception as e: print("  ERR", e)
PY
  tail -2 $o/${tag}_$name.err | cut -c1-300
}
timeout 300 python -m pytest tests/test_dist_gpu.py -x -q 2>&1 | tail -2
for c in 1 2 4; do GCNB_PULL_CTAS=$c runm 8 spmm_n8_pull_ctas$c --exchange pull; done
runm 4 spmm_n4_pull --exchange pull
runm 2 spmm_n2_pull --exchange pull
runm 8 epoch_n8_pull --exchange pull --mode epoch
runm 8 rmat_s24_n8_pull --exchange pull --workload rmat --scale 24 --edge-factor 16 --k 128

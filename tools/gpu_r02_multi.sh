#!/bin/bash
# round 2, multi-GPU batch (gpurun --gpus N): NCCL two-rank tests under the real barrier, bench lines with parity_check
# usage: tools/gpu_r02_multi.sh <tag> <N> [tests]
tag=${1:-r02_j}; N=${2:-2}; o=gpurun_out; mkdir -p $o
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
if [ "$3" = "tests" ]; then
  timeout 900 python -m pytest tests/test_dist_gpu.py tests/test_zz_wire_gpu.py tests/test_gemm_gpu.py -m gpu -x -q -k "two_ranks or wire or partitioned or second_device" > $o/${tag}_pytest_n$N.log 2>&1
  echo "pytest exit $?" >> $o/${tag}_pytest_n$N.log; tail -6 $o/${tag}_pytest_n$N.log
fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
IFS=";" read -ra MODES <<< "${MODES:-pull f32;allgather f32;pull bf16}"
for mode in "${MODES[@]}"; do
  set -- $mode
  f=$o/${tag}_spmm_n${N}_$1_$2.json
  timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 --exchange $1 --wire $2 > $f 2> ${f%.json}.err
  echo "spmm n=$N $1 $2 rc=$? $(grep -o '"ms_median": [0-9.]*\|"parity_check": {[^}]*}\|"ms_per_epoch": [0-9.]*' $f | cut -c1-160 | paste - - -)"
done
[ -n "$NO_REF" ] && exit 0
f=$o/${tag}_reference_n$N.json
timeout 600 $TR bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $f 2> ${f%.json}.err
echo "reference arm n=$N rc=$? $(grep -o '"cores": [0-9]*\|"value": [0-9.]*' $f | head -2 | paste - -)"

#!/bin/bash
# N-GPU pass (run under `gpurun --gpus N`): row-partitioned SpMM and GCN epoch at 1..N ranks.
tag=${1:-r01_m}; N=${2:-2}; o=gpurun_out; mkdir -p $o
shift 2
nvidia-smi topo -m > $o/${tag}_topo.txt 2>&1
for n in $(seq 2 $N); do
  case $n in 2|4|8) ;; *) continue;; esac
  for mode in spmm epoch; do
    NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
      --master-port $((29500 + n)) bench.py --gpus $n --mode $mode --steps 10 --warmup 3 "$@" \
      > $o/${tag}_${mode}_n$n.json 2> $o/${tag}_${mode}_n$n.err
    echo "$mode n=$n rc=$?"; cut -c1-400 $o/${tag}_${mode}_n$n.json; tail -3 $o/${tag}_${mode}_n$n.err
  done
done

#!/bin/bash
# round 2, multi-GPU batch ai: rendezvous through flags in peer memory (gcnb_peer_barrier) against the NCCL all-reduce;
# NCCL two-rank tests, timeline of the overlapped step, bench lines with the back-to-back parity check
tag=${1:-r02_ai}; N=${2:-2}; o=gpurun_out; mkdir -p $o
timeout 900 python -m pytest tests/test_dist_gpu.py tests/test_zz_wire_gpu.py -m gpu -x -q > $o/${tag}_pytest_n$N.log 2>&1
echo "pytest exit $? $(tail -n 1 $o/${tag}_pytest_n$N.log)"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
for e in "GCNB_FLAG_BARRIER=0" "X=1" "GCNB_PULL_TMA=32"; do
  env $e timeout 600 $TR tools/dist_timeline.py 2>>$o/${tag}_err.log | grep '^{' | tee -a $o/${tag}_timeline_n$N.jsonl | cut -c1-330
done
run() {
  name=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  f=$o/${tag}_n${N}_$name.json
  env "${envs[@]}" timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline "$@" > $f 2> ${f%.json}.err
  echo "$name rc=$? $(grep -o '"ms_per_step": [0-9.]*\|"ms_median": [0-9.]*\|"max_rel_err": [0-9.e-]*\|"ms_per_epoch": [0-9.]*' $f | cut -c1-200 | paste -s -d' ')"
}
run nccl_barrier GCNB_FLAG_BARRIER=0 -- --no-epoch --no-e2e
run flag_barrier X=1 --
run flag_barrier_tma GCNB_PULL_TMA=32 --
run allgather X=1 -- --exchange allgather --no-epoch --no-e2e

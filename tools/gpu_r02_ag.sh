#!/bin/bash
# round 2, multi-GPU batch ag (gpurun --gpus N): hub rows of small matrices cut into short pieces (GCNB_SHORT_PIECES),
# halo pull through the TMA unit (GCNB_PULL_TMA = ring depth) against the LDG pull kernel
# usage: tools/gpu_r02_ag.sh <tag> <N>
tag=${1:-r02_ag}; N=${2:-2}; o=gpurun_out; mkdir -p $o
GCNB_PULL_TMA=32 python -m pytest tests/test_dist_gpu.py -m gpu -x -q -k "single_process or two_ranks or partitioned_gcn" > $o/${tag}_pytest_tma.log 2>&1; echo "pytest tma rc=$? $(tail -n 1 $o/${tag}_pytest_tma.log)"
python tools/remote_block_probe.py --parts 8 --rank 3 2>>$o/${tag}_err.log | tail -n 1 | tee -a $o/${tag}_remote_block.jsonl
GCNB_SHORT_PIECES=0 python tools/remote_block_probe.py --parts 8 --rank 3 2>>$o/${tag}_err.log | tail -n 1 | tee -a $o/${tag}_remote_block.jsonl
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
run() {
  name=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  f=$o/${tag}_n${N}_$name.json
  env "${envs[@]}" timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --no-cpu-baseline "$@" > $f 2> ${f%.json}.err
  echo "$name rc=$? $(grep -o '"ms_per_step": [0-9.]*\|"ms_median": [0-9.]*\|"phase_ms_alone": {[^}]*}\|"max_rel_err": [0-9.e-]*\|"ms_per_epoch": [0-9.]*' $f | cut -c1-200 | paste -s -d' ')"
}
run base GCNB_SHORT_PIECES=0 -- --no-epoch
run short X=1 -- --no-epoch
run short_tma32 GCNB_PULL_TMA=32 -- --no-epoch
run short_tma16 GCNB_PULL_TMA=16 -- --no-epoch
run short_tma32_c2 GCNB_PULL_TMA=32 GCNB_PULL_CTAS=2 -- --no-epoch
run short_ldg_c2 GCNB_PULL_CTAS=2 -- --no-epoch
run short_tma32_epoch GCNB_PULL_TMA=32 --

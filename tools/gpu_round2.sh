#!/bin/bash
# reorder sweep, epoch lines, R-MAT points (1 GPU).  Usage: tools/gpu_round2.sh TAG
tag=${1:-r01_x}; o=gpurun_out; mkdir -p $o
run() { name=$1; shift; timeout ${TMO:-600} python bench.py "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo "$name rc=$?"; cut -c1-420 $o/${tag}_$name.json; tail -2 $o/${tag}_$name.err; }
Q="--no-cpu-baseline --no-e2e"
run products $Q
run products_bf16 $Q --dtype bf16
run reddit_bf16 $Q --workload reddit --dtype bf16
for ord in deg rcm gorder; do run arxiv_$ord $Q --workload arxiv --order $ord; done
for ord in deg rcm; do run reddit_$ord $Q --workload reddit --order $ord; done
run products_rcm $Q --order rcm
run products_shuffled $Q --shuffle
run products_shuffled_rcm $Q --shuffle --order rcm
for w in cora arxiv reddit products; do run epoch_$w --mode epoch --workload $w --steps 10 --warmup 3; done
run rmat_s22_ef16_k128 $Q --workload rmat --scale 22 --edge-factor 16 --k 128
run rmat_s22_ef16_k32 $Q --workload rmat --scale 22 --edge-factor 16 --k 32
run rmat_s22_ef64_k128 $Q --workload rmat --scale 22 --edge-factor 64 --k 128
TMO=900 run rmat_s24_ef16_k128 $Q --workload rmat --scale 24 --edge-factor 16 --k 128
TMO=900 run reddit_gorder $Q --workload reddit --order gorder

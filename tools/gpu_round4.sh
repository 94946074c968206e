#!/bin/bash
tag=${1:-r01_j}; o=gpurun_out; mkdir -p $o
run() { name=$1; shift; timeout ${TMO:-600} python bench.py "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo "$name rc=$?"; cut -c1-200 $o/${tag}_$name.json; tail -2 $o/${tag}_$name.err; }
timeout 900 python -m pytest tests/test_dist_gpu.py -x -q > $o/${tag}_pytest_dist.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest_dist.log
tail -15 $o/${tag}_pytest_dist.log
Q="--no-cpu-baseline --no-e2e --steps 10 --warmup 3"
run f32_pitch128 $Q --pitch 128
run bf16_pitch100 $Q --dtype bf16
run bf16_pitch104 $Q --dtype bf16 --pitch 104
run bf16_pitch128 $Q --dtype bf16 --pitch 128
run bf16_k104 $Q --dtype bf16 --k 104
run bf16_k128_p128 $Q --dtype bf16 --k 128
run f32_k128 $Q --k 128

#!/bin/bash
tag=${1:-r01_k}; o=gpurun_out; mkdir -p $o
run() { name=$1; shift; timeout ${TMO:-600} python bench.py "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo "$name rc=$?"; cut -c1-200 $o/${tag}_$name.json; tail -2 $o/${tag}_$name.err; }
timeout 900 python -m pytest tests/test_dist_gpu.py tests/test_spmm_gpu.py -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log
tail -15 $o/${tag}_pytest.log
Q="--no-cpu-baseline --no-e2e --steps 20 --warmup 5"
run bf16_k100 $Q --dtype bf16
run bf16_k128 $Q --dtype bf16 --k 128
run bf16_k256 $Q --dtype bf16 --k 256
run f32_k256 $Q --k 256
run reddit_bf16 $Q --dtype bf16 --workload reddit
run arxiv_bf16 $Q --dtype bf16 --workload arxiv

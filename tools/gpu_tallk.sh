#!/bin/bash
# one short GPU pass for the tall-K X·W variant: parity of both paths, their times at products scale,
# then the epoch line with it
tag=${1:-r01_ao}
mkdir -p gpurun_out
export PYTHONPATH=$PWD
timeout 60 python -m pytest tests/test_gemm_gpu.py tests/test_layers_gpu.py -m gpu -x -q > gpurun_out/${tag}_pytest_gemm_layers.log 2>&1
echo "pytest exit $?" >> gpurun_out/${tag}_pytest_gemm_layers.log
timeout 40 python tools/tallk_time.py > gpurun_out/${tag}_tallk.jsonl 2> gpurun_out/${tag}_tallk.err
timeout 60 python bench.py --mode epoch --workload products --steps 5 --warmup 3 > gpurun_out/${tag}_epoch_products.json 2> gpurun_out/${tag}_epoch_products.err
timeout 40 python -m pytest tests/test_dist_gpu.py -m gpu -x -q > gpurun_out/${tag}_pytest_dist.log 2>&1
echo "pytest exit $?" >> gpurun_out/${tag}_pytest_dist.log
tail -3 gpurun_out/${tag}_pytest_gemm_layers.log; cat gpurun_out/${tag}_tallk.jsonl; cat gpurun_out/${tag}_epoch_products.json; tail -2 gpurun_out/${tag}_pytest_dist.log

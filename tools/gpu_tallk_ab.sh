#!/bin/bash
# products-shaped epoch with the prepared-image dW (GCNB_XW_TALLK=0) and the tall-K variant, same box
tag=${1:-r01_ap}
mkdir -p gpurun_out
export PYTHONPATH=$PWD
GCNB_XW_TALLK=0 timeout 16 python bench.py --mode epoch --workload products --steps 5 --warmup 3 > gpurun_out/${tag}_epoch_products_image.json 2> /dev/null
GCNB_XW_TALLK=1 timeout 16 python bench.py --mode epoch --workload products --steps 5 --warmup 3 > gpurun_out/${tag}_epoch_products_tallk.json 2> /dev/null
GCNB_XW_MAX_CHAIN=128 timeout 16 python bench.py --mode epoch --workload products --steps 5 --warmup 3 > gpurun_out/${tag}_epoch_products_tallk_chain128.json 2> /dev/null
for f in gpurun_out/${tag}_epoch_products_*.json; do echo $f; cut -c1-120 $f; done

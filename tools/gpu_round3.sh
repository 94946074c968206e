#!/bin/bash
# 1-GPU pass: parity suite, headline + reference arm, epoch lines, reorder sweep (arxiv/products),
# R-MAT points, ncu of the epoch (launch list) and one full capture of the tcgen05 X·W kernel.
tag=${1:-r01_h}; o=gpurun_out; mkdir -p $o
run() { name=$1; shift; timeout ${TMO:-600} python bench.py "$@" > $o/${tag}_$name.json 2> $o/${tag}_$name.err; echo "$name rc=$?"; cut -c1-300 $o/${tag}_$name.json; tail -2 $o/${tag}_$name.err; }
timeout 1500 python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log
tail -3 $o/${tag}_pytest.log
run products
run reference --impl reference --steps 3 --warmup 1
Q="--no-cpu-baseline --no-e2e"
for w in cora arxiv reddit products; do run epoch_$w --mode epoch --workload $w --steps 10 --warmup 3; done
for ord in deg rcm gorder; do run arxiv_$ord $Q --workload arxiv --order $ord; done
run products_rcm $Q --order rcm
run products_shuffled $Q --shuffle
run rmat_s22_ef16_k128 $Q --workload rmat --scale 22 --edge-factor 16 --k 128
run rmat_s22_ef16_k32 $Q --workload rmat --scale 22 --edge-factor 16 --k 32
run rmat_s22_ef64_k128 $Q --workload rmat --scale 22 --edge-factor 64 --k 128
TMO=900 run rmat_s24_ef16_k128 $Q --workload rmat --scale 24 --edge-factor 16 --k 128
CMD="python bench.py --mode epoch --workload arxiv --steps 2 --warmup 1"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $o/${tag}_launches_epoch_arxiv.csv $CMD > $o/${tag}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:xw_kernel -s 2 -c 2 \
    -f -o $o/${tag}_prof_xw $CMD > $o/${tag}_ncu_xw.log 2>&1
ls -la $o | tail -30

#!/bin/bash
# One GPU-box pass: parity suite, bench lines of the BASELINE shapes, ncu launch list and one full
# capture of the dominant kernel.  Usage (from the repo root, under gpurun): tools/gpu_round.sh TAG
tag=${1:-r01_x}
o=gpurun_out
mkdir -p $o
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $o/${tag}_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log
tail -3 $o/${tag}_pytest.log
timeout 600 python bench.py > $o/${tag}_bench_products.json 2> $o/${tag}_bench_products.err; echo "bench exit $?"
cat $o/${tag}_bench_products.json
for w in arxiv reddit; do
  timeout 600 python bench.py --workload $w --no-e2e > $o/${tag}_bench_$w.json 2> $o/${tag}_bench_$w.err
  cat $o/${tag}_bench_$w.json
done
timeout 600 python bench.py --dtype bf16 --no-cpu-baseline > $o/${tag}_bench_products_bf16.json 2> $o/${tag}_bench_products_bf16.err
cat $o/${tag}_bench_products_bf16.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $o/${tag}_bench_reference.json 2> $o/${tag}_bench_reference.err
cat $o/${tag}_bench_reference.json
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 $CMD > $o/${tag}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv \
    --log-file $o/${tag}_launches.csv $CMD > $o/${tag}_ncu_launches.log 2>&1
timeout 600 $CMD > $o/${tag}_plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:spmm_ -s 3 -c 2 \
    -f -o $o/${tag}_prof $CMD > $o/${tag}_ncu_full.log 2>&1
ls -la $o | tail -20

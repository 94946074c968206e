#!/bin/bash
tag=${1:-r01_ah}; o=gpurun_out; mkdir -p $o
timeout 1500 python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -3 $o/${tag}_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > $o/${tag}_bench_default.json 2> $o/${tag}_bench_default.err; echo "bench rc=$?"; cut -c1-250 $o/${tag}_bench_default.json
timeout 300 python bench.py --impl reference --mode epoch --workload cora --steps 5 --warmup 2 > $o/${tag}_reference_epoch_cora.json 2>/dev/null; cut -c1-400 $o/${tag}_reference_epoch_cora.json
timeout 600 python bench.py --impl reference --mode epoch --workload arxiv --steps 3 --warmup 1 > $o/${tag}_reference_epoch_arxiv.json 2>/dev/null; cut -c1-400 $o/${tag}_reference_epoch_arxiv.json
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gcnb -c 100 --csv \
    --log-file $o/${tag}_launches_products_k100.csv $CMD > $o/${tag}_ncu_launches.log 2>&1; tail -1 $o/${tag}_ncu_launches.log | cut -c1-200

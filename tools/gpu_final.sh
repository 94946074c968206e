#!/bin/bash
tag=${1:-r01_z}; o=gpurun_out; mkdir -p $o
timeout 1500 python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -3 $o/${tag}_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $o/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $o/${tag}_smoke.log
timeout 900 python bench.py > $o/${tag}_bench_default.json 2> $o/${tag}_bench_default.err; echo "bench rc=$?"; cat $o/${tag}_bench_default.json; tail -2 $o/${tag}_bench_default.err
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $o/${tag}_bench_reference.json 2> $o/${tag}_bench_reference.err; echo "ref rc=$?"; cut -c1-300 $o/${tag}_bench_reference.json
timeout 600 python bench.py --order rabbit --workload arxiv --no-cpu-baseline --no-e2e > $o/${tag}_bench_arxiv_rabbit.json 2> $o/${tag}_bench_arxiv_rabbit.err; cut -c1-600 $o/${tag}_bench_arxiv_rabbit.json

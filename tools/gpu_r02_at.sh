#!/bin/bash
# round 2, GPU batch at: state of the round on one B200 — full GPU suite, smoke, default bench line, launch list of the
# bench, row pitch against the 128-byte line once more (the kernels now sit at 85-88 % of the LSU data pipe), epochs
tag=${1:-r02_at}; o=gpurun_out; mkdir -p $o
timeout 1500 python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -n 3 $o/${tag}_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1
timeout 900 python bench.py > $o/${tag}_bench_default.json 2> $o/${tag}_bench_default.err; echo "bench rc=$?"; cut -c1-330 $o/${tag}_bench_default.json
B="python bench.py --steps 20 --warmup 5 --comparators none --no-epoch --no-cpu-baseline --no-e2e"
for a in "100 100" "100 104" "100 128" "48 48" "48 64"; do set -- $a
  line=$($B --k $1 --pitch $2 2>>$o/${tag}_err.log | tail -n 1)
  echo "{\"k\": $1, \"pitch\": $2, \"line\": $line}" >> $o/${tag}_pitch.jsonl
  echo "k=$1 pitch=$2 $(echo "$line" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d.get("ms_median"), d.get("ms_min"))')"
done
for w in arxiv reddit cora; do
  python bench.py --mode epoch --workload $w --steps 10 --warmup 3 --no-cpu-baseline > $o/${tag}_epoch_$w.json 2>>$o/${tag}_err.log
  echo "epoch $w $(grep -o '"ms_per_step": [0-9.]*' $o/${tag}_epoch_$w.json)"
done
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --comparators none"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches_bench_default.csv $CMD > $o/${tag}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"

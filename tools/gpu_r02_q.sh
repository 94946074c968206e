#!/bin/bash
# round 2, GPU batch q: which stage bounds the streaming X·W kernel — GCNB_XS_DEBUG bits switch off one stage
# at a time (results wrong, times only); Reddit reorder sweep with bench-style timing; GPU suite for gemm/layers
tag=${1:-r02_q}; o=gpurun_out; mkdir -p $o
for dbg in 0 1 2 3 4 8 12 15; do
  GCNB_XS_DEBUG=$dbg XW_TIME_VARIANTS=1 XW_TIME_ONLY=plain timeout 200 python tools/xw_time.py 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:160]); continue
    d['dbg']=$dbg; print(json.dumps(d))" >> $o/${tag}_xs_debug.jsonl
done
python - <<PY
import json
for l in open("$o/${tag}_xs_debug.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"dbg {d['dbg']:2d} {d['ms']:8.3f} ms  {d['product'][:48]}")
PY
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_layers_gpu.py -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -3 $o/${tag}_pytest.log
timeout 300 python tools/xw_time.py > $o/${tag}_xw_shapes.jsonl 2>&1; python - <<PY
import json
for l in open("$o/${tag}_xw_shapes.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"{d['ms']:8.3f} ms {d['GBps']:8.1f} GB/s  {d['product'][:48]:48s}")
PY
f=$o/${tag}_epoch_products.json
timeout 600 python bench.py --mode epoch --workload products --steps 10 --warmup 3 > $f 2> ${f%.json}.err
echo "epoch products rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
timeout 1500 python tools/reddit_sweep.py > $o/${tag}_reddit_sweep.jsonl 2> $o/${tag}_reddit_sweep.err; echo "reddit sweep rc=$?"; python - <<PY
import json
for l in open("$o/${tag}_reddit_sweep.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    if "order" in d: print(f"{d['order']:9s} {d['class']:6s} k={d['k']:4d} {d['ms_median']:8.3f} ms  host {d['host_order_s']} s  chk {d['checksum_rel_err']:.1e}")
PY

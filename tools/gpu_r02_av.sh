#!/bin/bash
# round 2, GPU batch av: after making 36 warps/SM the fp32 stream default — SpMM / layer / dist suites, default bench line
tag=${1:-r02_av}; o=gpurun_out; mkdir -p $o
timeout 900 python -m pytest tests/test_spmm_gpu.py tests/test_layers_gpu.py tests/test_dist_gpu.py tests/test_zzz_no_alloc.py -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> $o/${tag}_pytest.log; tail -n 2 $o/${tag}_pytest.log
timeout 900 python bench.py > $o/${tag}_bench_default.json 2> $o/${tag}_bench_default.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("$o/${tag}_bench_default.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step", "ms_median", "ms_min")}, d["roofline"]["frac"], d["epoch"]["ms_per_epoch"], d["clocks"])
PY
python tools/remote_block_probe.py --parts 8 --rank 3 2>>$o/${tag}_err.log | tail -n 1 | tee -a $o/${tag}_remote_block.jsonl | cut -c1-400

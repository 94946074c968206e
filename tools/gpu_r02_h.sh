#!/bin/bash
# round 2, GPU batch h: ring-depth sensitivity of the streaming X·W kernel on the products-shaped epoch's products
tag=${1:-r02_h}; o=gpurun_out; mkdir -p $o
out=$o/${tag}_xw_shapes.jsonl; : > $out
timeout 300 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q 2>&1 | tail -2
for cfg in "GCNB_XS_LO=2" "GCNB_XS_LO=3" "GCNB_XS_LO=4" "GCNB_XS_LO=2 GCNB_XS_RA=3" "GCNB_XS_LO=3 GCNB_XW_WRES=0" "GCNB_XW_STREAM=0"; do
  env $cfg timeout 300 python tools/xw_time.py >> $out 2>&1
done
python - <<PY
import json
for l in open("$out"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"{d['ms']:8.3f} ms {d['GBps']:8.1f} GB/s  {d['product'][:40]:40s} {d['env']}")
PY

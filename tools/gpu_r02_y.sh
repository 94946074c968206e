#!/bin/bash
# round 2, GPU batch y: 128-column tiles with a 128 KB resident W image vs 64-column tiles (layer-1 product)
tag=${1:-r02_y}; o=gpurun_out; mkdir -p $o
for kb in 64 128; do
  GCNB_XS_WRES_KB=$kb XW_TIME_VARIANTS=1 XW_TIME_ONLY="fwd1" timeout 200 python tools/xw_time.py 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:160]); continue
    d['wres_kb']=$kb; print(json.dumps(d))" >> $o/${tag}_xw_wres_kb.jsonl
done
python - <<PY
import json
for l in open("$o/${tag}_xw_wres_kb.jsonl"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print(f"wres_kb {d['wres_kb']:4d} {d['ms']:8.3f} ms  {d['product'][:48]}")
PY

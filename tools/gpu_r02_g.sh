#!/bin/bash
# round 2, GPU batch g: Reddit epoch with the panel class opt-in, source-level stall profile of the streaming
# X·W kernel (summaries made on the box)
tag=${1:-r02_g}; o=gpurun_out; mkdir -p $o
f=$o/${tag}_epoch_reddit.json
timeout 600 python bench.py --mode epoch --workload reddit --steps 10 --warmup 3 > $f 2> ${f%.json}.err
echo "epoch reddit rc=$? $(grep -o '"ms_median": [0-9.]*\|"ms_min": [0-9.]*' $f | paste - -)"
CMD="python bench.py --mode epoch --steps 2 --warmup 2"
$CMD > /dev/null 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:xs_kernel -s 6 -c 3 -f -o /tmp/prof_xs $CMD > /tmp/ncu_xs.log 2>&1
echo "ncu xs rc=$?"
ncu -i /tmp/prof_xs.ncu-rep --page source --csv > /tmp/xs_source.csv 2>/dev/null
python - <<'PY' > gpurun_out/r02_g_ncu_xs_source_top.txt 2>&1
import csv
rows = list(csv.reader(open("/tmp/xs_source.csv")))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "data": []}; blocks.append(cur)
    elif cur is not None and cur["hdr"] is None: cur["hdr"] = r
    elif cur is not None and len(r) == len(cur["hdr"]): cur["data"].append(r)
for b in blocks:
    hdr, data = b["hdr"], b["data"]
    ia, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    tot = sum(int(r[isamp]) for r in data) or 1
    print("=====", b["name"][:100]); print("samples", tot, "warp-instructions", sum(int(r[iex]) for r in data), "SASS lines", len(data))
    sc = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    agg = {hdr[i]: sum(int(r[i] or 0) for r in data) for i in sc}
    print("stall totals:", ", ".join(f"{k[6:]} {100*v/tot:.1f}%" for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
    top = sorted(enumerate(data), key=lambda x: -int(x[1][isamp]))[:45]
    for idx, r in sorted(top):
        st = sorted(((hdr[i][6:], int(r[i] or 0)) for i in sc), key=lambda x: -x[1])[:2]
        print(f"{idx:5d} {r[ia].strip()[:70]:70s} {100*int(r[isamp])/tot:5.1f}% exec {r[iex]:>10s} {st}")
PY
ncu -i /tmp/prof_xs.ncu-rep --page details --csv 2>/dev/null | python -c "
import csv,sys
for r in csv.reader(sys.stdin):
    if len(r)>14 and any(k in r[12] for k in ('Duration','Throughput','Stall','Eligible','Issued','Active Warps','Registers','Hit Rate','Pipe')): print(r[4][:40],'|',r[11][:30],'|',r[12][:50],'|',r[13],'|',r[14])
" > gpurun_out/r02_g_ncu_xs_details.txt 2>&1
rm -f /tmp/prof_xs.ncu-rep; du -sh $o; head -60 gpurun_out/r02_g_ncu_xs_source_top.txt

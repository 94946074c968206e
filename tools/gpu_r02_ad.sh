#!/bin/bash
# round 2, GPU batch ad: fp32 feature rows through 256-bit loads (GCNB_F32_V8): parity of the new instantiations, then
# time against the 128-bit kernels on the products shape (k = 32..256) and the Reddit shape (k = 128, 602)
tag=${1:-r02_ad}; o=gpurun_out; mkdir -p $o
GCNB_F32_V8=3 python -m pytest tests/test_spmm_gpu.py -m gpu -x -q > $o/${tag}_pytest_v8.log 2>&1; echo "pytest v8 rc=$?"
GCNB_F32_V8=3 GCNB_PACKED_CFG=1 GCNB_STREAM_CFG=1 python -m pytest tests/test_spmm_gpu.py -m gpu -x -q -k "width or pitch or packed" > $o/${tag}_pytest_v8_cfg1.log 2>&1; echo "pytest v8 cfg1 rc=$?"
B="python bench.py --steps 20 --warmup 5 --comparators none --no-epoch --no-cpu-baseline --no-e2e"
run() {  # name, env..., -- args
  name=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  line=$(env "${envs[@]}" $B "$@" 2>>$o/${tag}_err.log | tail -n 1)
  echo "{\"run\": \"$name\", \"env\": \"${envs[*]}\", \"args\": \"$*\", \"line\": $line}" >> $o/${tag}_v8_sweep.jsonl
  echo "$name $(echo "$line" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print(d["ms_per_step"], d.get("ms_median"), d.get("ms_min"))' 2>/dev/null)"
}
run k100_p100_v4 X=1 -- --k 100
run k100_p104_v4 X=1 -- --k 100 --pitch 104
for c in 0 1 2 3 4; do run k100_p104_v8_cfg$c GCNB_F32_V8=3 GCNB_PACKED_CFG=$c -- --k 100 --pitch 104; done
run k128_v4 X=1 -- --k 128
for c in 0 1 2; do run k128_v8_cfg$c GCNB_F32_V8=3 GCNB_PACKED_CFG=$c -- --k 128; done
run k64_v4 X=1 -- --k 64
for c in 0 1 2; do run k64_v8_cfg$c GCNB_F32_V8=3 GCNB_PACKED_CFG=$c -- --k 64; done
run k32_v4 X=1 -- --k 32
for c in 0 1; do run k32_v8_cfg$c GCNB_F32_V8=3 GCNB_PACKED_CFG=$c -- --k 32; done
run k256_v4 X=1 -- --k 256
for c in 0 1; do run k256_v8_scfg$c GCNB_F32_V8=3 GCNB_STREAM_CFG=$c -- --k 256; done
run reddit_k128_v4 X=1 -- --workload reddit --k 128
for c in 0 1 2; do run reddit_k128_v8_cfg$c GCNB_F32_V8=3 GCNB_PACKED_CFG=$c -- --workload reddit --k 128; done
run reddit_k602_v4 X=1 -- --workload reddit --k 602 --pitch 608
for c in 0 1; do run reddit_k602_v8_scfg$c GCNB_F32_V8=3 GCNB_STREAM_CFG=$c -- --workload reddit --k 602 --pitch 608; done
tail -n 3 $o/${tag}_pytest_v8.log $o/${tag}_pytest_v8_cfg1.log

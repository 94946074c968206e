#!/bin/bash
# round 2, multi-GPU batch ay: pull kernel at 48 registers (it held 31 % of the register file beside the local block)
tag=${1:-r02_ay}; N=${2:-2}; o=gpurun_out; mkdir -p $o
timeout 600 python -m pytest tests/test_dist_gpu.py tests/test_zz_wire_gpu.py -m gpu -x -q > $o/${tag}_pytest_n$N.log 2>&1; echo "pytest exit $? $(tail -n 1 $o/${tag}_pytest_n$N.log)"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR tools/dist_timeline.py 2>>$o/${tag}_err.log | grep '^{' | tee -a $o/${tag}_timeline_n$N.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['rank'], 'step', round(d['step_ms'], 3), 'whole', round(d['whole_ms'], 3), 'bar', round(d['barrier_ms'], 3), 'pull', round(d['pull_ms'], 3), 'local', round(d['local_ms'], 3), 'remote', round(d['remote_ms'], 3), d['env'])"
f=$o/${tag}_spmm_n${N}_pull_f32.json
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > $f 2> ${f%.json}.err
echo "bench rc=$? $(grep -o '"ms_per_step": [0-9.]*\|"ms_median": [0-9.]*\|"phase_ms_alone": {[^}]*}\|"max_rel_err": [0-9.e-]*\|"ms_per_epoch": [0-9.]*' $f | cut -c1-220 | paste -s -d' ')"

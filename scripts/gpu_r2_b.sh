#!/bin/bash
# round-2 GPU pass B (2 GPUs): multi-process DP checks, all-reduce timeline, 2-GPU bench A/B
cd "$(dirname "$0")/.."
O=gpurun_out/r2b; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29551 scripts/gpu_dp_mp_check.py > $O/dp_mp_check.json 2> $O/dp_mp_check.err; echo "mpcheck rc=$?"
cat $O/dp_mp_check.json | head -c 3000; echo
timeout 200 $TR --master-port 29552 scripts/gpu_ar_timeline.py > $O/ar_timeline_2gpu.json 2> $O/ar_timeline.err; echo "timeline rc=$?"
cat $O/ar_timeline_2gpu.json | head -c 3000; echo
timeout 300 $TR --master-port 29553 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_2gpu_s20.json 2> $O/bench_2gpu_s20.err; echo "bench2 rc=$?"
CIB_EARLY_UPDATE=0 timeout 300 $TR --master-port 29554 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_2gpu_s20_noearly.json 2> $O/bench_2gpu_s20_noearly.err
for f in $O/bench_*.json; do echo "== $f"; python -c "
import json,sys
for l in open('$f'):
    if l.startswith('{'):
        d=json.loads(l); print({k:d.get(k) for k in ('value','ms_per_step','kernels_per_step','replicas_identical')}, d['e2e']['ms_per_step'])
"; done
timeout 600 python -m pytest tests/test_gpu_dist.py tests/test_gpu_engine.py -m gpu -x -q 2>&1 | tail -15

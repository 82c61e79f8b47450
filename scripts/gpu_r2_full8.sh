#!/bin/bash
# one 8-GPU session: bench at N=8 and N=1 (+ graphed-PyTorch baseline at 8), all-reduce timeline at 8, full-size workloads
cd "$(dirname "$0")/.."
O=gpurun_out/r2_full8; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29701 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench_ours_8gpu.json 2> $O/bench_ours_8gpu.err; echo "bench8 rc=$?"
timeout 120 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_ours_1gpu.json 2> $O/bench_ours_1gpu.err; echo "bench1 rc=$?"
timeout 200 $TR --master-port 29702 bench.py --impl torch-graph --gpus 8 --steps 20 --warmup 5 > $O/bench_torchgraph_8gpu.json 2> $O/bench_torchgraph_8gpu.err; echo "tg8 rc=$?"
timeout 100 $TR --master-port 29703 scripts/gpu_ar_timeline.py > $O/ar_timeline_8gpu.json 2> $O/ar_timeline_8gpu.err; echo "timeline rc=$?"
python - <<PY
import json
for n in ('bench_ours_8gpu','bench_ours_1gpu','bench_torchgraph_8gpu'):
    try:
        for l in open('$O/%s.json'%n):
            if l.startswith('{'):
                d=json.loads(l); print(n, {k:d.get(k) for k in ('value','ms_per_step','replicas_identical')}, (d.get('e2e') or {}).get('ms_per_step'))
    except Exception as e: print(n, e)
PY
bash scripts/gpu_r2_workloads.sh 8 full
cp gpurun_out/r2w_8/*.json $O/ 2>/dev/null

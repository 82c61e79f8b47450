#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out/r2_final8b; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29821 bench.py --gpus 8 --steps 20 --warmup 5 2>$O/b8.err | grep "^{" > $O/bench_ours_8gpu.json; echo "rc=$?"
timeout 120 python bench.py --gpus 1 --steps 20 --warmup 5 2>$O/b1.err | grep "^{" > $O/bench_ours_1gpu.json
python - <<PY
import json
for n in ('bench_ours_8gpu','bench_ours_1gpu'):
    d=json.loads(open('$O/%s.json'%n).readline()); print(n, d['value'], d['ms_per_step'], d.get('replicas_identical'), d['e2e']['value'], d['e2e']['ms_per_step'], d['clocks'])
PY

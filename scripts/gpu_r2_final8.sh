#!/bin/bash
# final 8-GPU confirmation: bench (default P2P push path, then NVLS), multi-process correctness at 8 ranks
cd "$(dirname "$0")/.."
O=gpurun_out/r2_final8; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29801 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench_ours_8gpu.json 2> $O/bench_ours_8gpu.err; echo "bench8 rc=$?"
CIB_NVLS=1 timeout 200 $TR --master-port 29802 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench_ours_8gpu_nvls.json 2> $O/bench_ours_8gpu_nvls.err; echo "bench8 nvls rc=$?"
timeout 120 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_ours_1gpu.json 2> $O/bench_ours_1gpu.err; echo "bench1 rc=$?"
timeout 300 $TR --master-port 29803 scripts/gpu_dp_mp_check.py > $O/dp_mp_check_8gpu.json 2> $O/dp_mp_check_8gpu.err; echo "mpcheck rc=$?"
python - <<PY
import json
for n in ('bench_ours_8gpu','bench_ours_8gpu_nvls','bench_ours_1gpu'):
    try:
        for l in open('$O/%s.json'%n):
            if l.startswith('{'):
                d=json.loads(l); print(n, {k:d.get(k) for k in ('value','ms_per_step','replicas_identical')}, (d.get('e2e') or {}).get('ms_per_step'), d['config'].get('allreduce'))
    except Exception as e: print(n, e)
for l in open('$O/dp_mp_check_8gpu.json'):
    if l.startswith('{'):
        d=json.loads(l); print('mp check 8:', d['ok'], {k:(v['ok'], v.get('max_rel_err')) for k,v in d['gradient'].items()}, {k:v['ok'] for k,v in d['replicas'].items()})
PY

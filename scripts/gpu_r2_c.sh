#!/bin/bash
# round-2 GPU pass C (N GPUs): all-reduce timeline + bench A/B switches
cd "$(dirname "$0")/.."
N=${1:-2}
O=gpurun_out/r2c_$N; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
P=29560
run() { name=$1; shift; P=$((P+1)); env "$@" timeout 300 $TR --master-port $P bench.py --gpus $N --steps 20 --warmup 5 > $O/bench_$name.json 2> $O/bench_$name.err; echo "$name rc=$?"; }
timeout 200 $TR --master-port 29559 scripts/gpu_ar_timeline.py > $O/ar_timeline.json 2> $O/ar_timeline.err; echo "timeline rc=$?"
run default X=1
run noearly CIB_EARLY_UPDATE=0
run nvls CIB_NVLS=1
run noll CIB_AR_LL=0
if [ "$2" = "more" ]; then
  run ctas16 CIB_PAR_CTAS=16
  run oneshot_all CIB_AR_MODE=1
  run pull CIB_AR_PULL=1 CIB_EARLY_UPDATE=0
fi
python - <<PY
import json,glob
for f in sorted(glob.glob('$O/bench_*.json')):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l); print(f.split('/')[-1], {k:d.get(k) for k in ('value','ms_per_step','kernels_per_step','replicas_identical')}, 'e2e', round(d['e2e']['ms_per_step'],4))
d=json.load(open('$O/ar_timeline.json'))
for r in d[:2]:
    print(r['rank'], r['steps'][-1])
PY
tail -q -n 3 $O/*.err | head -30

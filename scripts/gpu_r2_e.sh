#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out/r2e; mkdir -p $O
b() { name=$1; shift; env "$@" python bench.py --steps 100 --warmup 10 2>$O/$name.err | tee $O/bench_$name.json | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$name', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), 'loss', d['final_loss'])"; }
for v in "$@"; do b "$(echo $v | tr '=' '_')" $v; done

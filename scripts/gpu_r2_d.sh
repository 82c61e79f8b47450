#!/bin/bash
# 1-GPU A/B of the step-level switches + GPU test tier
cd "$(dirname "$0")/.."
O=gpurun_out/r2d; mkdir -p $O
b() { name=$1; shift; env "$@" python bench.py --steps 100 --warmup 10 2>$O/$name.err | tee $O/bench_$name.json | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$name', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), 'k/step', d['kernels_per_step'], 'loss', d['final_loss'])"; }
b default X=1
b serial_bwd CIB_SERIAL_BWD=1
b nofuse_act CIB_FUSE_DENSE_ACT=0
b nopdl CIB_PDL=0
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8

#!/bin/bash
cd "$(dirname "$0")/.."
O=gpurun_out/r2e; mkdir -p $O
b() { name=$1; shift; env "$@" python bench.py --steps 100 --warmup 10 2>$O/$name.err | tee $O/bench_$name.json | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$name', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],4), 'loss', d['final_loss'])"; }
b default X=1
b dgrad_bn64_noauto CIB_DGRAD_BN=64 CIB_GEMM_AUTOBN=0
b dgrad_bn32 CIB_DGRAD_BN=32
b dgrad_bn64 CIB_DGRAD_BN=64
b noauto CIB_GEMM_AUTOBN=0
timeout 600 python -m pytest tests/test_gpu_engine.py tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -4

#!/bin/bash
# Multi-GPU measurement pass (run on an N-GPU box through gpurun): correctness of the fused collectives, the all-reduce
# sweep against NCCL, and bench.py at 1/2/4/8 GPUs for both our engine and the stock-PyTorch baseline.
NG=${1:-8}
OUT=gpurun_out
mkdir -p $OUT
export NCCL_DEBUG=WARN
nvidia-smi topo -m > $OUT/topo_${NG}gpu.txt 2>&1
timeout 300 python scripts/gpu_dist_checks.py $NG > $OUT/dist_checks_${NG}gpu.log 2>&1; echo "dist_checks rc=$?"
PORT=29500
for n in 2 4 8; do
  [ $n -le $NG ] || continue
  PORT=$((PORT+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT \
    scripts/gpu_allreduce_sweep.py --out $OUT/allreduce_sweep_${n}gpu.json > $OUT/allreduce_sweep_${n}gpu.log 2>&1; echo "sweep $n rc=$?"
done
timeout 200 python bench.py --gpus 1 > $OUT/bench_ours_1gpu.json 2> $OUT/bench_ours_1gpu.err; echo "bench 1 rc=$?"
for n in 2 4 8; do
  [ $n -le $NG ] || continue
  for impl in ours torch; do
    PORT=$((PORT+1))
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $n --impl $impl > $OUT/bench_${impl}_${n}gpu.json 2> $OUT/bench_${impl}_${n}gpu.err; echo "bench $impl $n rc=$?"
  done
done
tail -n 3 $OUT/bench_*gpu.json

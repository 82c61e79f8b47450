#!/bin/bash
# round-2 GPU pass A: GPU test tier, 1-GPU bench (ours, torch-graph, torch eager), 2-GPU bench + multi-process checks
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/r2a
O=gpurun_out/r2a
nvidia-smi -L > $O/gpus.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_1gpu_s20.json 2> $O/bench_1gpu_s20.err; echo "bench1 rc=$?"
timeout 300 python bench.py --gpus 1 --steps 200 --warmup 20 > $O/bench_1gpu_s200.json 2> $O/bench_1gpu_s200.err
CIB_EARLY_UPDATE=0 timeout 300 python bench.py --gpus 1 --steps 200 --warmup 20 > $O/bench_1gpu_s200_noearly.json 2>&1
timeout 300 python bench.py --impl torch-graph --gpus 1 --steps 200 --warmup 20 > $O/bench_torchgraph_1gpu.json 2> $O/bench_torchgraph_1gpu.err; echo "tg rc=$?"
timeout 300 python bench.py --impl torch --gpus 1 --steps 200 --warmup 20 > $O/bench_torch_1gpu.json 2> $O/bench_torch_1gpu.err
NG=$(nvidia-smi -L | wc -l)
if [ "$NG" -ge 2 ]; then
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
  timeout 300 $TR --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_2gpu_s20.json 2> $O/bench_2gpu_s20.err; echo "bench2 rc=$?"
  CIB_EARLY_UPDATE=0 timeout 300 $TR --master-port 29542 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_2gpu_s20_noearly.json 2> $O/bench_2gpu_s20_noearly.err
  CIB_AR_PULL=1 CIB_EARLY_UPDATE=0 timeout 300 $TR --master-port 29543 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_2gpu_s20_pull.json 2> $O/bench_2gpu_s20_pull.err
  timeout 300 $TR --master-port 29544 bench.py --impl torch-graph --gpus 2 --steps 20 --warmup 5 > $O/bench_torchgraph_2gpu.json 2> $O/bench_torchgraph_2gpu.err; echo "tg2 rc=$?"
fi
for f in $O/bench_*.json; do echo "== $f"; head -c 1500 $f | tail -c 1500; echo; done
tail -3 $O/*.err | tail -40

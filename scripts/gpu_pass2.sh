#!/bin/bash
# 1-GPU pass: HPO packing exploration
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_engine.py -x -q -m gpu -k "concurrent or partial" 2>&1 | tail -15
for s in 2 4 8; do
  timeout 120 python scripts/hpo_bench.py --workload mnist --trials 16 --engines 1 --slots $s --out $OUT/hpo_mnist_1gpu_s$s.json > $OUT/hpo_mnist_1gpu_s$s.log 2>&1; echo "hpo mnist slots=$s rc=$?"; tail -n 1 $OUT/hpo_mnist_1gpu_s$s.log | cut -c1-600
done
timeout 120 python scripts/hpo_bench.py --workload rpv --trials 4 --engines 1 --slots 2 --out $OUT/hpo_rpv_1gpu_s2.json > $OUT/hpo_rpv_1gpu_s2.log 2>&1; echo "hpo rpv rc=$?"; tail -n 1 $OUT/hpo_rpv_1gpu_s2.log | cut -c1-600
mkdir -p $OUT/cl; cp ~/.cib200/clusters/*.log $OUT/cl/ 2>/dev/null

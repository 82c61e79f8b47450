#!/bin/bash
# 8-GPU HPO throughput: BASELINE.json configs 3 and 4 (+ config 1, the CPU MNIST epoch).
OUT=gpurun_out; mkdir -p $OUT
timeout 200 python scripts/hpo_bench.py --workload mnist --trials 64 --slots 4 --out $OUT/hpo_mnist_8gpu_s4.json > $OUT/hpo_mnist_8gpu_s4.log 2>&1; echo "mnist s4 rc=$?"; tail -n 1 $OUT/hpo_mnist_8gpu_s4.log | cut -c1-700
timeout 200 python scripts/hpo_bench.py --workload mnist --trials 64 --slots 1 --out $OUT/hpo_mnist_8gpu_s1.json > $OUT/hpo_mnist_8gpu_s1.log 2>&1; echo "mnist s1 rc=$?"; tail -n 1 $OUT/hpo_mnist_8gpu_s1.log | cut -c1-700
timeout 300 python scripts/hpo_bench.py --workload rpv --trials 32 --slots 2 --widgets --out $OUT/hpo_rpv_8gpu_widgets.json > $OUT/hpo_rpv_8gpu_widgets.log 2>&1; echo "rpv widgets rc=$?"; tail -n 1 $OUT/hpo_rpv_8gpu_widgets.log | cut -c1-900
timeout 120 python scripts/cpu_mnist_baseline.py 2>&1 | tail -n 1
timeout 200 python examples/dist_train_rpv.py --engines 8 --quick --device gpu > $OUT/example_dist_train_rpv_8gpu.log 2>&1; echo "dist_train_rpv example rc=$?"; tail -n 6 $OUT/example_dist_train_rpv_8gpu.log

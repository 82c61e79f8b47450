#!/bin/bash
# The reference's real workloads on N GPUs of one box, each recorded as JSON (wall time, clocks, metrics) under gpurun_out/r2w_N/.
#   gpu_r2_workloads.sh N full   : notebook-size configurations (DistTrain_rpv, DistTrain_mnist, CrayHPO_rpv 128 evaluations, HPO random searches)
#   gpu_r2_workloads.sh N quick  : the same code paths at reduced sizes (validation run)
cd "$(dirname "$0")/.."
N=${1:-2}; MODE=${2:-quick}
O=gpurun_out/r2w_$N; mkdir -p $O
export SCRATCH=/tmp/cib_scratch; mkdir -p $SCRATCH
export RPV_CHANNELS=3 CIB_SYNTH_LABEL_NOISE=0.05
PER=$(( N >= 8 ? 4 : (N >= 2 ? 2 : 1) ))
if [ "$MODE" = "full" ]; then POP=8; DEMES=4; GENS=4; MT=64; RT=32; EP=16; SC=1.0; else POP=2; DEMES=1; GENS=2; MT=8; RT=4; EP=2; SC=0.05; fi
t() { /usr/bin/time -f "%e s wall" "$@"; }
echo "== genetic HPO (cluster launcher, $((N/PER)) concurrent x $PER-GPU data-parallel evaluations)"
timeout 1500 python examples/cray_hpo_rpv.py --device gpu --nodes $N --nodes-per-eval $PER --launcher cluster --pop-size $POP --num-demes $DEMES --generations $GENS \
   --channels 3 --out-dir $SCRATCH --record $O/cray_hpo_rpv_cluster.json > $O/cray_hpo_rpv_cluster.log 2>&1; echo "rc=$?"; tail -2 $O/cray_hpo_rpv_cluster.log | cut -c1-300
if [ "$MODE" != "full" ]; then
  echo "== genetic HPO (one torchrun job per evaluation: the reference's way)"
  timeout 900 python examples/cray_hpo_rpv.py --device gpu --nodes $N --nodes-per-eval $PER --launcher wlm --pop-size 2 --num-demes 1 --generations 1 \
     --channels 3 --out-dir $SCRATCH --record $O/cray_hpo_rpv_torchrun.json > $O/cray_hpo_rpv_torchrun.log 2>&1; echo "rc=$?"; tail -1 $O/cray_hpo_rpv_torchrun.log | cut -c1-300
fi
echo "== DistTrain_rpv"
QUICK=""; [ "$MODE" = "full" ] || QUICK=""
timeout 900 python examples/dist_train_rpv.py --device gpu --engines $N --out-dir $SCRATCH --record $O/dist_train_rpv.json > $O/dist_train_rpv.log 2>&1; echo "rc=$?"; tail -3 $O/dist_train_rpv.log | cut -c1-300
echo "== DistTrain_mnist"
timeout 900 python examples/dist_train_mnist.py --device gpu --engines $N --out-dir $SCRATCH --record $O/dist_train_mnist.json > $O/dist_train_mnist.log 2>&1; echo "rc=$?"; tail -2 $O/dist_train_mnist.log | cut -c1-300
echo "== HPO random search, MNIST ($MT trials, 4 per GPU)"
timeout 1200 python scripts/hpo_bench.py --workload mnist --trials $MT --epochs $EP --scale $SC --slots 4 --out $O/hpo_mnist.json > $O/hpo_mnist.log 2>&1; echo "rc=$?"; tail -1 $O/hpo_mnist.log | cut -c1-400
echo "== HPO random search, RPV 64x64x3 with the live widget ($RT trials, 2 per GPU)"
timeout 1200 python scripts/hpo_bench.py --workload rpv --widgets --trials $RT --epochs $EP --scale $SC --slots 2 --out $O/hpo_rpv_widgets.json > $O/hpo_rpv_widgets.log 2>&1; echo "rc=$?"; tail -1 $O/hpo_rpv_widgets.log | cut -c1-400
python - <<PY
import json,glob
for f in sorted(glob.glob('$O/*.json')):
    d=json.load(open(f))
    keep={k:d[k] for k in ('wall_s','train_wall_s','train_samples_per_s_all_ranks','evaluations','evaluations_per_hour','eval_secs_mean','finite_foms','best_fom','trials_per_hour','best_val_acc','completed','checkpoint_files','auc') if k in d}
    print(f.split('/')[-1], keep)
PY

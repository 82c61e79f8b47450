#!/bin/bash
# Batch data-parallel RPV training on N GPUs of this box (reference: sbatch + srun -n $SLURM_NNODES python train_rpv.py).
#   NGPUS=8 ./batch_scripts/train_rpv.sh --n-epochs 4 --lr-scaling linear
cd "$(dirname "$0")/.."
. ./setup.sh
NGPUS="${NGPUS:-$(python -c 'import torch; print(max(torch.cuda.device_count(), 1))')}"
mkdir -p batch_logs
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NGPUS" --master-addr 127.0.0.1 --master-port "${MASTER_PORT:-29533}" \
    train_rpv.py "$@" 2>&1 | tee "batch_logs/train_rpv-$$.out"

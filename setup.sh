# Environment for interactive deep learning on one 8xB200 box (source it: `. setup.sh`).
# Counterpart of the reference setup.sh (Cori `module load` + thread variables): here there are no modules; we only
# make the package importable, build the sm_100a library if it is missing, and set the knobs the framework reads.
export CIB_ROOT="$(cd "$(dirname "${BASH_SOURCE[0]:-$0}")" && pwd)"
export PYTHONPATH="$CIB_ROOT${PYTHONPATH:+:$PYTHONPATH}"
[ -f "$CIB_ROOT/cori_intml_examples_b200/lib/libcib200.so" ] || "$CIB_ROOT/scripts/build.sh"
# where cluster connection files and engine logs go (the reference used $SCRATCH)
export CIB_CLUSTER_DIR="${CIB_CLUSTER_DIR:-${SCRATCH:-$HOME}/.cib200/clusters}"
# CPU-side threads for the oracle backend / data preparation (reference: NUM_INTER_THREADS / NUM_INTRA_THREADS)
export NUM_INTER_THREADS="${NUM_INTER_THREADS:-2}"
export NUM_INTRA_THREADS="${NUM_INTRA_THREADS:-8}"
export OMP_NUM_THREADS="$NUM_INTRA_THREADS"

#!/bin/bash
# compute-sanitizer targets: memcheck / racecheck / synccheck / initcheck over the stand-alone kernel cases, plus a
# short engine run.  Usage on a GPU box:  bash scripts/sanitize.sh [tool] [case-filter]     (default: memcheck, a fast subset)
# Host side: SANITIZE=address|thread bash scripts/build.sh adds -fsanitize=... to the host compiler flags of the library.
TOOL=${1:-memcheck}
FILTER=${2:-"gemm_kk_k32 conv_fwd_16_32_valid dgrad_16_16_valid wgrad_16_32_valid conv1_fwd_1_32_valid conv1_wgrad_1_32_valid dense_64_2304"}
OUT=gpurun_out/sanitizer; mkdir -p $OUT
rc_all=0
for c in $FILTER; do
  timeout 600 compute-sanitizer --tool $TOOL --error-exitcode 77 --log-file $OUT/${TOOL}_$c.log \
      python scripts/gpu_kernel_checks.py --case $c > $OUT/${TOOL}_$c.out 2>&1
  rc=$?
  errs=$(grep -c "ERROR SUMMARY: 0 errors" $OUT/${TOOL}_$c.log)
  echo "$TOOL $c rc=$rc clean=$errs"
  [ "$errs" = "1" ] || rc_all=1
done
exit $rc_all

#!/bin/bash
# ncu --set full captures (one GPU): the tcgen05 kernels of one RPV training step and of one legacy-model step
cd "$(dirname "$0")/.."
O=gpurun_out/ncu_r2; mkdir -p $O
CH=3 STEPS=3 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"conv_tc2_kernel|wgrad_tc2_kernel|conv_first|gemm_tc_kernel|head_mc" -s 30 -c 14 -o $O/rpv_step -f python scripts/gpu_step_profile.py rpv > $O/rpv_step.log 2>&1; echo "rc=$?"
STEPS=2 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"conv_tc_kernel|wgrad_tc_kernel" -s 15 -c 8 -o $O/legacy_step -f python scripts/gpu_step_profile.py legacy > $O/legacy_step.log 2>&1; echo "rc=$?"
ls -la $O

#!/bin/bash
# P.V (gemm_tc_fused_kernel) with different cache hints on the streamed P rows
mkdir -p gpurun_out; : > gpurun_out/r01d_pv_hint.log
for h in 0 1 2 3; do
  echo -n "hint $h: " >> gpurun_out/r01d_pv_hint.log
  WF2CU_FUSED_AHINT=$h timeout 60 python scripts/attn_profile.py >> gpurun_out/r01d_pv_hint.log 2>&1
done
cat gpurun_out/r01d_pv_hint.log

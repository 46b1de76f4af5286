#!/bin/bash
# P.V (gemm_tc_fused_kernel) with different L2 bulk-prefetch settings for the P rows: "blocks,ahead" in 128-byte k-blocks
mkdir -p gpurun_out; : > gpurun_out/r01d_pv_prefetch.log
for pf in 0,0 4,8 8,8 8,16 16,16 28,32; do
  echo -n "prefetch $pf: " >> gpurun_out/r01d_pv_prefetch.log
  WF2CU_FUSED_PREFETCH=$pf timeout 60 python scripts/attn_profile.py >> gpurun_out/r01d_pv_prefetch.log 2>&1
done
cat gpurun_out/r01d_pv_prefetch.log

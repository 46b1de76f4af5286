#!/bin/bash
mkdir -p gpurun_out
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc_fused_kernel" -s 2 -c 1 -o gpurun_out/attn_pv_prof python scripts/attn_profile.py > gpurun_out/attn_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/attn_pv_prof.ncu-rep --page raw --csv > gpurun_out/attn_pv_prof_raw.csv 2>/dev/null
ls -la gpurun_out/attn_pv_prof*; tail -n 3 gpurun_out/attn_ncu.log

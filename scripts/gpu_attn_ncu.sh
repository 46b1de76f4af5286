#!/bin/bash
# Full ncu capture of GPT-2's P.V launch (gemm_tc_fused_kernel, 96 x (1024 x 1024 x 64)).
mkdir -p gpurun_out
timeout 60 python scripts/attn_profile.py > gpurun_out/r01d_attn.log 2>&1 || exit 1
timeout 200 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_fused_kernel -s 2 -c 1 -o gpurun_out/attn_pv_prof python scripts/attn_profile.py > gpurun_out/attn_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/attn_pv_prof.ncu-rep --page raw --csv > gpurun_out/r01d_attn_pv_raw.csv 2>/dev/null
ncu -i gpurun_out/attn_pv_prof.ncu-rep --page source --csv > gpurun_out/r01d_attn_pv_source.csv 2>gpurun_out/attn_src.err
rm -f gpurun_out/attn_pv_prof.ncu-rep
ls -la gpurun_out/r01d_attn*; tail -n 3 gpurun_out/attn_ncu.log

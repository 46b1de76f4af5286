# GPT-2 (C5) at HEAD: attention products timed alone, then the launch list of two eager forwards.
mkdir -p gpurun_out
timeout 60 python scripts/attn_profile.py > gpurun_out/r01d_attn.log 2>&1; cat gpurun_out/r01d_attn.log
timeout 60 python scripts/gpt2_profile.py > gpurun_out/r01d_gpt2_plain.log 2>&1 && \
timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r01d_gpt2_launches.csv python scripts/gpt2_profile.py > gpurun_out/r01d_gpt2_ncu.log 2>&1
tail -n 2 gpurun_out/r01d_gpt2_plain.log gpurun_out/r01d_gpt2_ncu.log

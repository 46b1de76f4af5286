# Round-1 (d) evidence run on one B200 at the final state: tests, both bench arms, launch lists, full captures of the GEMM kernels.
set -x
mkdir -p gpurun_out
export_rep() { ncu -i gpurun_out/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null; rm -f gpurun_out/$1.ncu-rep; }
timeout 120 python -m pytest tests -m gpu -x -q > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/d_pytest.log; tail -n 3 gpurun_out/d_pytest.log
timeout 150 python bench.py > gpurun_out/d_bench.json 2> gpurun_out/d_bench.err; echo "bench rc=$?"
timeout 100 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/d_bench_ref.json 2> gpurun_out/d_bench_ref.err; echo "ref rc=$?"
# launch list of the sweep (the same command exited 0 inside the bench above)
timeout 100 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/d_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu --mlp 0 --gemm 0 --gpt2 0 --cnn 0 > gpurun_out/d_ncu1.log 2>&1
# GPT-2 launch list
timeout 100 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/d_gpt2_launches.csv python scripts/gpt2_profile.py > gpurun_out/d_gpt2_ncu.log 2>&1
# GEMM N=4096 (planes + persistent tcgen05 kernel with the bulk-store epilogue) and Q.K^T
timeout 100 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|split_planes" -s 6 -c 3 -o gpurun_out/d_gemm_prof python scripts/gemm_profile.py 4096 > gpurun_out/d_gemm_ncu.log 2>&1
export_rep d_gemm_prof
timeout 100 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc_kernel" -s 4 -c 1 -o gpurun_out/d_qk_prof python scripts/attn_profile.py > gpurun_out/d_qk_ncu.log 2>&1
export_rep d_qk_prof
du -sh gpurun_out; tail -n 2 gpurun_out/d_gemm_ncu.log gpurun_out/d_qk_ncu.log

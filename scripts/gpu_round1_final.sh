# Round-1 evidence run on one B200: tests, both bench arms, ncu launch lists and full captures.
# gpurun brings back at most 64 MiB: every .ncu-rep is exported to CSV on the box and deleted (the GEMM one, 3 kernels, is kept).
set -x
export_rep() { ncu -i gpurun_out/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null; rm -f gpurun_out/$1.ncu-rep; }
python -m pytest tests -m gpu -x -q > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest.log; tail -n 3 gpurun_out/f_pytest.log
python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/f_bench_ref.json 2> gpurun_out/f_bench_ref.err; echo "ref rc=$?"
# launch list of the sweep (the program exits 0 without ncu first)
python bench.py --steps 3 --warmup 3 --no-cpu --mlp 0 --gemm 0 --gpt2 0 > gpurun_out/f_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/f_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu --mlp 0 --gemm 0 --gpt2 0 > gpurun_out/f_ncu1.log 2>&1
# full capture of one sweep step (12 launches)
ncu --set full --clock-control none -s 36 -c 12 -o gpurun_out/f_prof python bench.py --steps 3 --warmup 3 --no-cpu --mlp 0 --gemm 0 --gpt2 0 > gpurun_out/f_ncu2.log 2>&1
export_rep f_prof
# MLP step launch list + full capture of its last step (15 launches)
python scripts/mlp_profile.py > gpurun_out/f_mlp_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/f_mlp_launches.csv python scripts/mlp_profile.py > gpurun_out/f_mlp_ncu.log 2>&1
ncu --set full --clock-control none -s 105 -c 15 -o gpurun_out/f_mlp_prof python scripts/mlp_profile.py > gpurun_out/f_mlp_ncu2.log 2>&1
export_rep f_mlp_prof
# GEMM N=4096 (planes + persistent tcgen05 kernel)
python scripts/gemm_profile.py 4096 > gpurun_out/f_gemm_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|split_planes" -s 6 -c 3 -o gpurun_out/f_gemm_prof python scripts/gemm_profile.py 4096 > gpurun_out/f_gemm_ncu.log 2>&1
ncu -i gpurun_out/f_gemm_prof.ncu-rep --page raw --csv > gpurun_out/f_gemm_prof_raw.csv 2>/dev/null
du -sh gpurun_out
tail -n 2 gpurun_out/f_ncu2.log gpurun_out/f_mlp_ncu2.log gpurun_out/f_gemm_ncu.log
python - <<'PY'
import json
d=json.loads(open('gpurun_out/f_bench.json').read().strip().splitlines()[-1])
print('value', d['value'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'], 'clocks', d['clocks'])
for k,v in d['cases'].items(): print(f"{k:18s} {v['ms']*1e3:7.2f} us {v['gbps']:7.0f} GB/s {v['frac']:.3f}")
print(d['mlp']['weak']); print(d['gpt2']); print([(c['N'], round(c['tflops'],1), round(c['pipe_frac'],3)) for c in d['gemm']['cases']])
print(d['cpu_baseline'])
print(open('gpurun_out/f_bench_ref.json').read()[:600])
PY

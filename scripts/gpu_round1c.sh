set -x
python bench.py > gpurun_out/c_bench_n1.json 2> gpurun_out/c_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/c_bench_n1.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/c_bench_n1.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value']); print(d['gpt2']); print(d['gemm']['roofline']); print(d['mlp']['weak'])
PY
python scripts/gemm_profile.py 4096 > gpurun_out/c_gemm_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|split_planes" -s 6 -c 3 -o gpurun_out/c_gemm_prof python scripts/gemm_profile.py 4096 > gpurun_out/c_gemm_ncu.log 2>&1
tail -4 gpurun_out/c_gemm_ncu.log

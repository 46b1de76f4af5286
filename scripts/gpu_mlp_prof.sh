python -m pytest tests -m gpu -x -q > gpurun_out/q_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 30 gpurun_out/q_pytest.log
python scripts/mlp_profile.py > gpurun_out/mlp_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/mlp_launches.csv python scripts/mlp_profile.py > gpurun_out/mlp_ncu.log 2>&1
tail -n 2 gpurun_out/mlp_plain.log gpurun_out/mlp_ncu.log
python bench.py --no-cpu --gemm 0 --gpt2 0 > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/q_bench.json').read().strip().splitlines()[-1])
print('value', d['value']); print(d['mlp']['weak'])
PY

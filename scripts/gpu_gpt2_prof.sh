python -m pytest tests -m gpu -x -q > gpurun_out/q_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 15 gpurun_out/q_pytest.log
python scripts/gpt2_profile.py > gpurun_out/gpt2_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/gpt2_launches.csv python scripts/gpt2_profile.py > gpurun_out/gpt2_ncu.log 2>&1
tail -n 2 gpurun_out/gpt2_plain.log gpurun_out/gpt2_ncu.log
python bench.py --no-cpu --gemm 0 --mlp 0 > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/q_bench.json').read().strip().splitlines()[-1])
print(d['gpt2'])"

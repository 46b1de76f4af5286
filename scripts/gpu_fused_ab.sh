timeout 200 python scripts/gemm_small.py > gpurun_out/gemm_small6.log 2>&1; echo rc=$?; head -8 gpurun_out/gemm_small6.log | cut -c1-150; grep failures gpurun_out/gemm_small6.log
python bench.py --no-cpu --gemm 0 --mlp-steps 100 > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/q_bench.json').read().strip().splitlines()[-1])
print(d['gpt2']['ms_per_forward'], d['mlp']['weak']['ms_per_step'])"

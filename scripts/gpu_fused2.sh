#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpt2.py tests/test_gpu_mlp.py -m gpu -x -q -k "matmul or gemm or gpt2 or attention or mlp or linear" 2>&1 | tail -n 4
timeout 200 python scripts/attn_profile.py
timeout 200 python scripts/gemm_small.py 2>&1 | tail -n 12
timeout 300 python bench.py --steps 5 --warmup 3 --gemm 0 --cnn 0 --no-cpu > gpurun_out/epi_bench.json 2> gpurun_out/epi_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/epi_bench.json').read().strip().splitlines()[-1])
print('gpt2', d.get('gpt2'))
print('mlp', d['mlp']['weak'])
PY

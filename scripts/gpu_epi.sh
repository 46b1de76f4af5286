#!/bin/bash
# coalesced epilogue + 256x192 tiles: parity, A/B timing, GPT-2 forward
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpt2.py -m gpu -x -q -k "matmul or gemm or gpt2 or attention" 2>&1 | tail -n 6
timeout 300 python scripts/gemm_t256.py > gpurun_out/gemm_t256.log 2>&1; echo "t256 rc=$?"; cat gpurun_out/gemm_t256.log
timeout 300 python bench.py --steps 5 --warmup 3 --mlp 0 --gemm 0 --cnn 0 --no-cpu > gpurun_out/epi_bench.json 2> gpurun_out/epi_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/epi_bench.json').read().strip().splitlines()[-1])
print('gpt2', d.get('gpt2'))
PY

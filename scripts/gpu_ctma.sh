#!/bin/bash
# bulk-tensor-store epilogue of the persistent tcgen05 kernels: tests, Q.K^T with it off / on, bench
mkdir -p gpurun_out
timeout 170 python -m pytest tests -m gpu -x -q > gpurun_out/ctma_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/ctma_pytest.log
echo -n "ctma off: "; WF2CU_GEMM_CTMA=0 timeout 60 python scripts/attn_profile.py
echo -n "ctma on : "; timeout 60 python scripts/attn_profile.py
timeout 170 python bench.py --no-cpu --mlp 0 > gpurun_out/ctma_bench.json 2> gpurun_out/ctma_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/ctma_bench.json').read().strip().splitlines()[-1])
print([(c['N'], round(c['ms']*1e3,1), round(c['tflops'],1)) for c in d['gemm']['cases']]); print(d['gpt2']['ms_per_forward'])
PY

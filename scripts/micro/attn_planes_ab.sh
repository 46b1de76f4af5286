WF2CU_ATTN_PLANES=1 python scripts/attn_bench.py > gpurun_out/r02g_attn_planes.log 2>&1; tail -2 gpurun_out/r02g_attn_planes.log
python scripts/attn_bench.py > gpurun_out/r02g_attn_raw.log 2>&1; tail -2 gpurun_out/r02g_attn_raw.log
WF2CU_ATTN_PLANES=1 python -m pytest tests/test_gpt2.py -m gpu -q 2>&1 | tail -3

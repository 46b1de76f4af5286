for d in 0 1 2 4 6 7; do echo "WF2CU_ATTN_DBG=$d"; WF2CU_ATTN_DBG=$d timeout 120 python scripts/attn_bench.py 2>&1 | tail -1; done
echo "planes"; for d in 0 7; do echo "WF2CU_ATTN_DBG=$d PLANES"; WF2CU_ATTN_PLANES=1 WF2CU_ATTN_DBG=$d timeout 120 python scripts/attn_bench.py 2>&1 | tail -1; done

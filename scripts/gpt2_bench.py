#!/usr/bin/env python
"""GPT-2 small forward (B = 8, N = 1024) as bench.py times it, alone: ms per forward under the CUDA graph, with and without
the fused attention kernels and the raw-operand GEMM."""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
wf = importlib.import_module("cl-waffe2_b200")
g = importlib.import_module("cl-waffe2_b200.gpt2")
wf.init(0)
for fuse in ([True, False] if os.environ.get("AB", "1") == "1" else [True]):
    g.FUSE_ATTENTION = fuse
    r = bench.gpt2_forward_bench(wf, oracle_check=fuse and os.environ.get("CHECK", "0") == "1")
    print(json.dumps({"fuse_attention": fuse, "ms_per_forward": r["ms_per_forward"], "launches": r["launches_per_forward"],
                      "max_rel_err": r.get("max_rel_err")}))

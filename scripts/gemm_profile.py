#!/usr/bin/env python
"""One MatmulNode (N x N fp32, tcgen05 3xTF32 path) a few times — the program ncu wraps."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
wf = importlib.import_module("cl-waffe2_b200")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
wf.init(0)
rng = np.random.default_rng(0)
A = wf.from_numpy(rng.uniform(-1, 1, size=(N, N)).astype(np.float32))
B = wf.from_numpy(rng.uniform(-1, 1, size=(N, N)).astype(np.float32))
o = wf.make_tensor((N, N))
for _ in range(3):
    wf.ops.matmul(A, B, out=o)
wf.ops.sync()
print("ok", float(o.to_numpy()[0, 0]))

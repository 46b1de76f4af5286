#!/usr/bin/env python
"""The element-wise sweep's dominant case, `(!add a b)` on (4096 4096) fp32 (benchmarks/arithmetic.lisp:11-24): copy + in-place
add fused into one three-operand launch — the program ncu wraps."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
wf = importlib.import_module("cl-waffe2_b200")
wf.init(0)
rng = np.random.default_rng(0)
a = wf.from_numpy(rng.normal(size=(4096, 4096)).astype(np.float32))
b = wf.from_numpy(rng.normal(size=(4096, 4096)).astype(np.float32))
out = wf.make_tensor((4096, 4096))
for _ in range(4):
    wf.ops.add(a, b, out=out)
    wf.ops.unary("sin", a, out=out)
    wf.ops.sum(a)
wf.ops.sync()
print("ok", float(out.to_numpy()[0, 0]))

#!/usr/bin/env python
"""One compiled MNIST-MLP training step (C4) replayed a few times — the program ncu wraps."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
wf = importlib.import_module("cl-waffe2_b200")
nn = importlib.import_module("cl-waffe2_b200.nn")
wf.init(0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
DIMS = [784, 256, 256, 10]
rng = np.random.default_rng(0)
X = rng.uniform(0, 1, size=(B, DIMS[0])).astype(np.float32)
Y = np.eye(10, dtype=np.float32)[rng.integers(10, size=B)]
model = nn.MLPSequence(DIMS, "relu", seed=1234)
step = nn.CompiledStep(model, "softmax-cross-entropy-sum", nn.Adam(model.arena, lr=1e-3), X.shape, Y.shape, world=1)
step.set_inputs(X, Y)
for _ in range(8):
    step.step()
wf.ops.sync()
print("ok loss", step.loss_value())

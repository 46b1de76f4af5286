#!/usr/bin/env python
"""One GPT-2 small forward (C5: B=8, N=1024, fp32) run eagerly twice — the program ncu wraps for a launch list."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
wf = importlib.import_module("cl-waffe2_b200")
g = importlib.import_module("cl-waffe2_b200.gpt2")
wf.init(0)
cfg = dict(g.CONFIG_SMALL)
model = g.GPT2(cfg, seed=0)
rng = np.random.default_rng(11)
ctx = wf.from_numpy(rng.integers(0, cfg["n_vocab"], size=(8, 1024)).astype(np.float32))
pool = wf.StaticPool()
for _ in range(2):
    with pool:
        out = model(ctx)
wf.ops.sync()
print("ok", out.shape)

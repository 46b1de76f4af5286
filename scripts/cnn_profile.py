"""MNIST-CNN training step: time per step and launches per step (scripts/gpu_conv.sh)."""
import argparse, importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=100)
ap.add_argument("--batch", type=int, default=0)
a = ap.parse_args()
wf = importlib.import_module("cl-waffe2_b200")
wf.init(0)
mb = importlib.import_module("cl-waffe2_b200.mlp_bench")
for b in ([a.batch] if a.batch else [100, 1024]):
    r = mb._run_one(None, 0, 1, b, a.steps, 5, workload="cnn")
    print(json.dumps({k: r[k] for k in ("per_gpu_batch", "ms_per_step", "samples_per_s", "e2e_ms_per_step", "launches_per_step", "loss_after")}))

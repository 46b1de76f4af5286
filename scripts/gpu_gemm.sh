set -x
timeout 600 python scripts/gemm_check.py --json gpurun_out/gemm_check.json > gpurun_out/gemm_check.log 2>&1; echo "gemm_check rc=$?"
tail -40 gpurun_out/gemm_check.log
timeout 200 python - <<'PY' > gpurun_out/torch_gemm_peaks.log 2>&1
import torch, time
for dt, name, tf32 in ((torch.float32, "tf32", True), (torch.float32, "fp32", False), (torch.bfloat16, "bf16", False)):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    for N in (4096, 8192):
        a = torch.randn(N, N, device="cuda", dtype=dt); b = torch.randn(N, N, device="cuda", dtype=dt)
        for _ in range(3): a @ b
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        best = 1e9
        for _ in range(5):
            e0.record(); c = a @ b; e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1))
        print(f"cublas {name} N={N}: {best:.4f} ms {2*N**3/best/1e9:.1f} TFLOP/s", flush=True)
PY
cat gpurun_out/torch_gemm_peaks.log

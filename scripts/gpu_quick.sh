# quick check: parity tests + the C2 sweep only (no cpu baseline, no mlp/gemm/gpt2)
python -m pytest tests -m gpu -x -q > gpurun_out/q_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/q_pytest.log
python bench.py --no-cpu --mlp 0 --gemm 0 --gpt2 0 > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/q_bench.json').read().strip().splitlines()[-1])
print('value', d['value'])
for k,v in d['cases'].items(): print(f"{k:18s} {v['ms']*1e3:7.2f} us {v['gbps']:7.0f} GB/s {v['frac']:.3f}")
PY

# usage: bash scripts/gpu_multi.sh <ngpus>
N=${1:-2}
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/m_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/m_pytest.log
python bench.py > gpurun_out/m_bench_n1.json 2> gpurun_out/m_bench_n1.err; echo "bench1 rc=$?"
for n in 2 4 8; do
  if [ $n -le $N ]; then
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n > gpurun_out/m_bench_n$n.json 2> gpurun_out/m_bench_n$n.err; echo "bench$n rc=$?"
    tail -5 gpurun_out/m_bench_n$n.err
  fi
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/m_bench_n*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d['value'], d.get('mlp',{}).get('weak'), d.get('mlp',{}).get('strong'))
        if d.get('gemm'): print([ (c['N'], round(c['tflops'],1)) for c in d['gemm']['cases']])
    except Exception as e: print(f, 'ERR', e)
PY

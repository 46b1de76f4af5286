# usage: bash scripts/gpu_scale.sh <ngpus>   (run under gpurun --gpus N)
N=${1:-2}
export WF2_BENCH_TRACE=1 WF2_BENCH_TRACE_AFTER=100
for n in 1 2 4 8; do
  if [ $n -le $N ]; then
    if [ $n -eq 1 ]; then
      timeout 300 python bench.py --no-cpu --gemm 0 --gpt2 0 > gpurun_out/sc_bench_n$n.json 2> gpurun_out/sc_bench_n$n.err; echo "bench$n rc=$?"
    else
      timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --no-cpu --gemm 0 --gpt2 0 > gpurun_out/sc_bench_n$n.json 2> gpurun_out/sc_bench_n$n.err; echo "bench$n rc=$?"
    fi
  fi
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/sc_bench_n*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        w=d['mlp']['weak']; s=d['mlp'].get('strong')
        print(f, 'sweep', round(d['value']), 'weak ms', round(w['ms_per_step'],4), 'samples/s', round(w['samples_per_s']), 'e2e', round(w['e2e_samples_per_s']), 'strong', s and (round(s['ms_per_step'],4), round(s['samples_per_s'])))
    except Exception as e: print(f, 'ERR', e)
PY

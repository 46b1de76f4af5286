set -x
export WF2_BENCH_TRACE=1 WF2_BENCH_TRACE_AFTER=100
N=${1:-2}
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/s_bench_n$N.json 2> gpurun_out/s_bench_n$N.err; echo "bench$N rc=$?"
grep -v "^W1104\|^\*\*\*\|OMP_NUM" gpurun_out/s_bench_n$N.err | tail -30
python - <<PY
import json
d=json.loads(open('gpurun_out/s_bench_n$N.json').read().strip().splitlines()[-1])
print(d['value'], d['n_gpus']); print(json.dumps(d['mlp'],indent=1))
PY

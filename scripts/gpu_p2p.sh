N=${1:-2}
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 scripts/p2p_check.py > gpurun_out/p2p_check_n$N.log 2>&1; echo "p2p_check rc=$?"
grep -v "^W1104\|^\*\*\*\|OMP_NUM" gpurun_out/p2p_check_n$N.log | tail -20
export WF2_BENCH_TRACE=1 WF2_BENCH_TRACE_AFTER=100
if [ -z "$SKIP_BENCH" ]; then timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-cpu --gemm 0 --gpt2 0 > gpurun_out/p2p_bench_n$N.json 2> gpurun_out/p2p_bench_n$N.err; echo "bench$N rc=$?"; fi
[ -z "$SKIP_BENCH" ] && python - <<PY
import json
d=json.loads(open('gpurun_out/p2p_bench_n$N.json').read().strip().splitlines()[-1])
m=d['mlp']
for k in ('weak','strong','weak_nccl'):
    if k in m: print(k, {kk: m[k][kk] for kk in ('exchange','ms_per_step','samples_per_s','replicas_identical','p2p_timed_out','launches_per_step')})
PY
grep -v "^W1104\|^\*\*\*\|OMP_NUM\|^\[" gpurun_out/p2p_bench_n$N.err | tail -5

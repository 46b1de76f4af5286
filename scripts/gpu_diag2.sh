set -x
python -m pytest tests/test_gpt2.py -m gpu -x -q > gpurun_out/d_gpt2.log 2>&1; echo "gpt2 rc=$?"; tail -15 gpurun_out/d_gpt2.log
export WF2_BENCH_TRACE=1 WF2_BENCH_TRACE_AFTER=60 NCCL_DEBUG=WARN
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --mlp-steps 20 --mlp-warmup 3 > gpurun_out/d_bench_n2.json 2> gpurun_out/d_bench_n2.err; echo "bench2 rc=$?"
grep -v "^W1104\|^\*\*\*\|OMP_NUM" gpurun_out/d_bench_n2.err | head -120
cat gpurun_out/d_bench_n2.json | head -c 1500

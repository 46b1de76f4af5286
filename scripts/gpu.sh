#!/bin/bash
# One runner for every GPU call of this repo (replaces the per-call gpu_*.sh scripts of round 1).
#   gpurun --timeout 900 -- 'bash scripts/gpu.sh <tag> <task> [task ...]'
# tasks:  tests            pytest -m gpu (all), log -> gpurun_out/<tag>_pytest.log
#         newtests         only tests/test_gpu_bench_shapes.py (prints kept)
#         bench            python bench.py            -> gpurun_out/<tag>_bench.json / .err
#         benchref         python bench.py --impl reference
#         benchnopdl       the sweep / fused chains / MLP with WF2CU_PDL=0 (A/B of programmatic dependent launch)
#         launches         ncu launch list of `bench.py --steps 2 --warmup 1 --gemm 0 --gpt2 0 --cnn 0 --mlp 0 --fused 1 --no-cpu`
#         ncu:<regex>:<script.py>   ncu --set full on kernels matching <regex> while running python <script.py>
#         py:<script.py>   run a python script, log -> gpurun_out/<tag>_<script>.log
# Every task runs under its own timeout so that a protocol bug cannot hold the box.
tag=$1; shift
mkdir -p gpurun_out
for task in "$@"; do
  case "$task" in
    tests)    timeout 1500 python -m pytest tests -m gpu -q -rf 2>&1 | tail -60 > gpurun_out/${tag}_pytest.log; tail -5 gpurun_out/${tag}_pytest.log ;;
    newtests) timeout 900 python -m pytest tests/test_gpu_bench_shapes.py -m gpu -q -rf -s 2>&1 | tail -80 > gpurun_out/${tag}_newtests.log; tail -40 gpurun_out/${tag}_newtests.log ;;
    bench)    timeout 900 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/${tag}_bench.json; tail -5 gpurun_out/${tag}_bench.err ;;
    benchnopdl) WF2CU_PDL=0 timeout 900 python bench.py --gemm 0 --gpt2 0 --cnn 0 --no-cpu > gpurun_out/${tag}_bench_nopdl.json 2> gpurun_out/${tag}_bench_nopdl.err; echo "bench (WF2CU_PDL=0) rc=$?" ;;
    benchref) timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "benchref rc=$?" ;;
    launches) cmd="python bench.py --steps 2 --warmup 1 --gemm 0 --gpt2 0 --cnn 0 --mlp 0 --no-cpu"
              timeout 300 $cmd > gpurun_out/${tag}_launches_plain.log 2>&1 &&
              timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv $cmd > gpurun_out/${tag}_launches_ncu.log 2>&1
              echo "launches rc=$?" ;;
    ncu:*)    IFS=: read -r _ regex script <<< "$task"; base=$(basename "$script" .py)
              timeout 300 python "$script" > gpurun_out/${tag}_${base}_plain.log 2>&1 &&
              timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$regex" -c 3 -f -o gpurun_out/${tag}_${base} python "$script" > gpurun_out/${tag}_${base}_ncu.log 2>&1
              echo "ncu $regex rc=$?"; tail -3 gpurun_out/${tag}_${base}_plain.log ;;
    py:*)     script=${task#py:}; base=$(basename "$script" .py)
              timeout 900 python $script > gpurun_out/${tag}_${base}.log 2>&1; echo "$script rc=$?"; tail -30 gpurun_out/${tag}_${base}.log ;;
    *) echo "unknown task $task" ;;
  esac
done

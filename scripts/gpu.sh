#!/bin/bash
# One runner for every GPU call of this repo (replaces the per-call gpu_*.sh scripts of round 1).
#   gpurun --timeout 900 -- 'bash scripts/gpu.sh <tag> <task> [task ...]'
# tasks:  tests            pytest -m gpu (all), log -> gpurun_out/<tag>_pytest.log
#         newtests         only tests/test_gpu_bench_shapes.py (prints kept)
#         pytest:<file>    one test file, prints kept
#         bench            python bench.py            -> gpurun_out/<tag>_bench.json / .err
#         benchref         python bench.py --impl reference
#         benchN:<n>       torchrun bench.py on n GPUs (call gpurun with --gpus n)
#         mlpN:<n>         the sharded MLP step only on n GPUs: LL (push) exchange, then WF2CU_P2P_LL=0 (pull), + scripts/p2p_check.py
#         benchnopdl       the sweep / fused chains / MLP with WF2CU_PDL=0 (A/B of programmatic dependent launch)
#         launches         ncu launch list of `bench.py --steps 2 --warmup 1 --gemm 0 --gpt2 0 --cnn 0 --mlp 0 --fused 1 --no-cpu`
#         ncu:<regex>:<script.py>[+arg...]   ncu --set full on kernels matching <regex> while running python <script.py> [args]
#         ab:<script.py>[+arg...]   the script under the built library and under every cl-waffe2_b200/lib/variants/*.so, in one call
#         ncutime:<script.py>  per-kernel device times of a script (ncu gpu__time_duration, cold-cache, serialised)
#         py:<script.py>   run a python script, log -> gpurun_out/<tag>_<script>.log
# Every task runs under its own timeout so that a protocol bug cannot hold the box.
tag=$1; shift
mkdir -p gpurun_out
for task in "$@"; do
  case "$task" in
    tests)    timeout 1500 python -m pytest tests -m gpu -q -rf 2>&1 | tail -60 > gpurun_out/${tag}_pytest.log; tail -5 gpurun_out/${tag}_pytest.log ;;
    pytest:*) f=${task#pytest:}; base=$(basename "$f" .py)
              timeout 900 python -m pytest "$f" -m gpu -q -rf -s 2>&1 | tail -60 > gpurun_out/${tag}_${base}.log; tail -25 gpurun_out/${tag}_${base}.log ;;
    newtests) timeout 900 python -m pytest tests/test_gpu_bench_shapes.py -m gpu -q -rf -s 2>&1 | tail -80 > gpurun_out/${tag}_newtests.log; tail -40 gpurun_out/${tag}_newtests.log ;;
    bench)    timeout 900 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/${tag}_bench.json; tail -5 gpurun_out/${tag}_bench.err ;;
    benchnopdl) WF2CU_PDL=0 timeout 900 python bench.py --gemm 0 --gpt2 0 --cnn 0 --no-cpu > gpurun_out/${tag}_bench_nopdl.json 2> gpurun_out/${tag}_bench_nopdl.err; echo "bench (WF2CU_PDL=0) rc=$?" ;;
    benchN:*) n=${task#benchN:}
              timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $n --steps 10 --warmup 3 \
                > gpurun_out/${tag}_bench_n${n}.json 2> gpurun_out/${tag}_bench_n${n}.err; echo "bench N=$n rc=$?"; tail -c 1400 gpurun_out/${tag}_bench_n${n}.json; grep -v "^W\|^\*\*\*" gpurun_out/${tag}_bench_n${n}.err | tail -8 ;;
    mlpN:*)   n=${task#mlpN:}
              for ll in 1 0; do
                WF2CU_P2P_TRACE=1 WF2CU_P2P_LL=$ll timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$ll bench.py --gpus $n --steps 5 --warmup 3 \
                  --gemm 0 --gpt2 0 --cnn 0 --fused 0 --cublas 0 > gpurun_out/${tag}_mlp_n${n}_ll${ll}.json 2> gpurun_out/${tag}_mlp_n${n}_ll${ll}.err
                echo "mlp N=$n LL=$ll rc=$?"; python -c "import json,sys; d=json.loads(open('gpurun_out/${tag}_mlp_n${n}_ll${ll}.json').read().strip().splitlines()[-1]); print(json.dumps(d['mlp']))"
                grep "p2p trace rank 0" gpurun_out/${tag}_mlp_n${n}_ll${ll}.err | tail -6
              done
              timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 scripts/p2p_check.py 2>&1 | grep -v "^W\|^\*\*\*" | tail -$((n + 2)) | tee gpurun_out/${tag}_p2p_check_n${n}.log ;;
    flat)     for v in "" "WF2CU_FLAT_STAGES=3" "WF2CU_FLAT_STAGES=3 WF2CU_FLAT_PDL=2" "WF2CU_FLAT_STAGES=2" "WF2CU_FLAT_STAGES=2 WF2CU_FLAT_PDL=2" "WF2CU_FLAT_STAGES=3 WF2CU_FLAT_PDL=0" "WF2CU_FLAT_STAGES=2 WF2CU_FLAT_CTAS=3" "WF2CU_FLAT_STAGES=3"; do env $v timeout 120 python scripts/flat_variants.py 2>&1 | tail -1; done | tee gpurun_out/${tag}_flat_variants.log ;;
    ab:*)     script=${task#ab:}; base=$(basename "${script%%+*}" .py)   # every lib/variants/*.so in turn (scripts/build_variant.sh), twice, same box
              cp cl-waffe2_b200/lib/libwaffe2cuda.so /tmp/libwaffe2cuda_base.so
              for rep in 1 2; do for v in /tmp/libwaffe2cuda_base.so cl-waffe2_b200/lib/variants/*.so; do
                cp $v cl-waffe2_b200/lib/libwaffe2cuda.so; echo "== $(basename $v) rep $rep"; timeout 300 python ${script//+/ } 2>&1 | tail -${AB_TAIL:-2}
              done; done | tee gpurun_out/${tag}_ab_${base}.log
              cp /tmp/libwaffe2cuda_base.so cl-waffe2_b200/lib/libwaffe2cuda.so ;;
    benchref) timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "benchref rc=$?" ;;
    launches) cmd="python bench.py --steps 2 --warmup 1 --gemm 0 --gpt2 0 --cnn 0 --mlp 0 --no-cpu"
              timeout 300 $cmd > gpurun_out/${tag}_launches_plain.log 2>&1 &&
              timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv $cmd > gpurun_out/${tag}_launches_ncu.log 2>&1
              echo "launches rc=$?" ;;
    ncu:*)    IFS=: read -r _ regex script <<< "$task"; base=$(basename "${script%%+*}" .py)$(echo "${script#*.py}" | tr -d '+')_$(echo "$regex" | tr -cd 'A-Za-z0-9_')
              cmd="python ${script//+/ }"          # scripts/x.py+8192 -> python scripts/x.py 8192
              timeout 300 $cmd > gpurun_out/${tag}_${base}_plain.log 2>&1 &&
              timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$regex" -c ${NCU_COUNT:-3} -f -o gpurun_out/${tag}_${base} $cmd > gpurun_out/${tag}_${base}_ncu.log 2>&1
              echo "ncu $regex rc=$?"; tail -3 gpurun_out/${tag}_${base}_plain.log ;;
    ncutime:*) script=${task#ncutime:}; base=$(basename "$script" .py)
              timeout 300 python $script > gpurun_out/${tag}_${base}_plain.log 2>&1 &&
              timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_${base}_launches.csv python $script > gpurun_out/${tag}_${base}_ncu.log 2>&1
              echo "ncutime $script rc=$?"
              python - <<PYEOF
import csv, collections
rows = [l for l in open("gpurun_out/${tag}_${base}_launches.csv") if not l.startswith("==")]
agg = collections.OrderedDict()
for r in csv.DictReader(rows):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        t = float(r["Metric Value"]); t = t / 1000 if r["Metric Unit"] in ("ns", "nsecond") else t
        a = agg.setdefault(r["Kernel Name"][:90], [0, 0.0]); a[0] += 1; a[1] += t
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
    print(f"{t / c:9.1f} us x {c:4d}  {k}")
PYEOF
              ;;
    py:*)     script=${task#py:}; base=$(basename "$script" .py)
              timeout 900 python $script > gpurun_out/${tag}_${base}.log 2>&1; echo "$script rc=$?"; tail -30 gpurun_out/${tag}_${base}.log ;;
    *) echo "unknown task $task" ;;
  esac
done

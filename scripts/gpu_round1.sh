set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1_pytest.log
python bench.py > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1_bench_ref.json 2> gpurun_out/r1_bench_ref.err; echo "ref rc=$?"
python bench.py --steps 3 --warmup 3 --no-cpu --mlp 0 > gpurun_out/r1_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu --mlp 0 > gpurun_out/r1_ncu1.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu --mlp 0 > gpurun_out/r1_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -s 40 -c 16 -o gpurun_out/r1_prof python bench.py --steps 3 --warmup 3 --no-cpu --mlp 0 > gpurun_out/r1_ncu2.log 2>&1
tail -3 gpurun_out/r1_pytest.log; cat gpurun_out/r1_bench.json | head -c 3000

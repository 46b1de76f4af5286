timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/h_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/h_smoke.log
( time timeout 900 python bench.py > gpurun_out/h_bench.json 2> gpurun_out/h_bench.err ) 2> gpurun_out/h_bench.time; echo "bench rc=$?"; tail -n 3 gpurun_out/h_bench.time
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/cnn1024_launches.csv python scripts/cnn_profile.py --steps 2 --batch 1024 > gpurun_out/cnn1024_ncu.log 2>&1; echo "ncu rc=$?"

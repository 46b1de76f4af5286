# usage: bash scripts/gpu_conv.sh   (under gpurun, 1 GPU): conv parity tests + CNN step timing + launch list
timeout 900 python -m pytest tests/test_gpu_conv.py -x -q > gpurun_out/conv_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 15 gpurun_out/conv_pytest.log
timeout 300 python scripts/cnn_profile.py > gpurun_out/cnn_plain.log 2>&1; echo "cnn rc=$?"; tail -n 12 gpurun_out/cnn_plain.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/cnn_launches.csv python scripts/cnn_profile.py --steps 3 --batch 100 > gpurun_out/cnn_ncu.log 2>&1; echo "ncu rc=$?"

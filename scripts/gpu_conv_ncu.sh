# full ncu capture of the skinny GEMM kernels inside one CNN step (batch 100)
timeout 600 ncu --set full --clock-control none -k regex:skinny -c 8 -o gpurun_out/cnn_skinny python scripts/cnn_profile.py --steps 1 --batch 100 > gpurun_out/cnn_skinny_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/cnn_skinny.ncu-rep --page raw --csv > gpurun_out/cnn_skinny_raw.csv 2>/dev/null; rm -f gpurun_out/cnn_skinny.ncu-rep; ls -la gpurun_out/cnn_skinny_raw.csv

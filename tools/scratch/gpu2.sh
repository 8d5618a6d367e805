mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python tools/prof_passes.py f32 1024 3 > gpurun_out/passes_f32.log 2>&1; cat gpurun_out/passes_f32.log
python tools/prof_passes.py f32 1024 1 > gpurun_out/plain2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:nb2_kernel -s 14 -c 20 -o gpurun_out/prof_r1a python tools/prof_passes.py f32 1024 1 > gpurun_out/ncu3.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu3.log; ls -la gpurun_out/*.ncu-rep

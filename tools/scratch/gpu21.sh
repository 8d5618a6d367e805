mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "l2_blocked" > gpurun_out/pytest21.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest21.log
timeout 900 python tools/prof_l2.py f32 1024 5 0,4,8,16,32,64 2>&1 | tail -8 | tee gpurun_out/l2_21.log
timeout 600 python tools/prof_l2.py f32 512 10 0,8,16,32,64,128 2>&1 | tail -8 | tee -a gpurun_out/l2_21.log

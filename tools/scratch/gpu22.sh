mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "l2_blocked" > gpurun_out/pytest22.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest22.log
timeout 900 python tools/prof_l2.py f32 1024 5 0,4x2,8x2,16x2,4x3,8x3,2x4,4x4,8x4 2>&1 | tail -10 | tee gpurun_out/l2_22.log
timeout 600 python tools/prof_l2.py f32 512 10 0,16x2,32x2,16x3,32x3,16x4 2>&1 | tail -8 | tee -a gpurun_out/l2_22.log

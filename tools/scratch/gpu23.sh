mkdir -p gpurun_out
timeout 900 python tools/prof_l2.py f32 1024 5 0,4x2,8x2,16x2,4x3,8x3,2x4,4x4,8x4 > gpurun_out/l2_23.log 2>&1; tail -12 gpurun_out/l2_23.log | cut -c1-180
timeout 600 python tools/prof_l2.py f32 512 10 0,16x2,32x2,16x3,64x2,16x4 > gpurun_out/l2_23b.log 2>&1; tail -8 gpurun_out/l2_23b.log | cut -c1-180

mkdir -p gpurun_out
CUDA_LAUNCH_BLOCKING=1 timeout 300 python tools/scratch/dbg_l2.py 128 2,4,8 > gpurun_out/dbg24.log 2>&1; grep -n "planes\|^ok\|Error" gpurun_out/dbg24.log | cut -c1-200
NB2_L2_STREAMS_X=1 timeout 300 python tools/scratch/dbg_l2.py 128 8,4,2,16 > gpurun_out/dbg24b.log 2>&1; grep -n "planes\|^ok\|Error" gpurun_out/dbg24b.log | cut -c1-200

mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_distributed.py -x -q > gpurun_out/pytest43.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest43.log | cut -c1-200
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench43_n2.json 2> gpurun_out/bench43_n2.err; echo "bench rc=$?"; cut -c1-230 gpurun_out/bench43_n2.json
NIFTYB200_FUSED_AXPBY=0 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench43_n2_unfused.json 2>/dev/null; echo "unfused rc=$?"; cut -c1-230 gpurun_out/bench43_n2_unfused.json

mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_distributed.py -x -q > gpurun_out/pytest32.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest32.log | cut -c1-250
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench32_n2.json 2> gpurun_out/bench32_n2.err; echo "bench rc=$?"; cut -c1-1200 gpurun_out/bench32_n2.json; tail -3 gpurun_out/bench32_n2.err

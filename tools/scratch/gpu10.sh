mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_distributed.py -q -x > gpurun_out/pytest_dist.log 2>&1; echo "pytest dist rc=$?"; tail -5 gpurun_out/pytest_dist.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench1 rc=$?"; wc -l gpurun_out/bench_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench2 rc=$?"; cat gpurun_out/bench_n2.json

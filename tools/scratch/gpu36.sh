mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_distributed.py -x -q > gpurun_out/pytest36.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest36.log | cut -c1-200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench36_n2.json 2> gpurun_out/bench36_n2.err; echo "bench rc=$?"; cut -c1-260 gpurun_out/bench36_n2.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/prof_slab.py > gpurun_out/slab36_n2.log 2>&1; tail -1 gpurun_out/slab36_n2.log | cut -c1-600

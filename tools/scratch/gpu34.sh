mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "strided_scatter or tma" > gpurun_out/pytest34a.log 2>&1; echo "pytest-a rc=$?"; tail -2 gpurun_out/pytest34a.log | cut -c1-200
timeout 900 python -m pytest tests/test_gpu_distributed.py -x -q > gpurun_out/pytest34.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest34.log | cut -c1-250
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench34_n2.json 2> gpurun_out/bench34_n2.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench34_n2.json; tail -2 gpurun_out/bench34_n2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/prof_slab.py > gpurun_out/slab34_n2.log 2>&1; tail -2 gpurun_out/slab34_n2.log | cut -c1-600

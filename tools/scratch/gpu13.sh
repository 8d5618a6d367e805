mkdir -p gpurun_out
N=${NG:-2}
timeout 600 python -m pytest tests/test_gpu_distributed.py -q -x > gpurun_out/pytest_dist.log 2>&1; echo "pytest dist rc=$?"; tail -3 gpurun_out/pytest_dist.log
NIFTYB200_SLAB_P2P=0 timeout 600 python -m pytest tests/test_gpu_distributed.py -q -x > gpurun_out/pytest_dist_nccl.log 2>&1; echo "pytest dist (nccl path) rc=$?"; tail -3 gpurun_out/pytest_dist_nccl.log
for cfg in "1 -" "1 4"; do
set -- $cfg
if [ "$2" = "-" ]; then unset NB2_STRIDED_LOGW; else export NB2_STRIDED_LOGW=$2; fi
NIFTYB200_SLAB_P2P=$1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29620 tools/prof_slab.py 1024 f32 2>&1 | grep "world=\|Error\|error" | tee -a gpurun_out/prof_slab.log
done
unset NB2_STRIDED_LOGW
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n${N}_p2p.json 2> gpurun_out/bench_n${N}_p2p.err; echo "bench rc=$?"; cut -c 1-400 gpurun_out/bench_n${N}_p2p.json

mkdir -p gpurun_out
N=${NG:-4}
for cfg in "1 0" "4 0" "4 1" "8 1" "2 1"; do
set -- $cfg
NIFTYB200_SLAB_CHUNKS=$1 TORCH_NCCL_HIGH_PRIORITY=$2 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29620 tools/prof_slab.py 1024 f32 2>&1 | grep "world=" | tee -a gpurun_out/prof_slab.log
done

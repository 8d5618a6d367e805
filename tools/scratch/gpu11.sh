mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_distributed.py -q -x > gpurun_out/pytest_dist.log 2>&1; echo "pytest dist rc=$?"; tail -3 gpurun_out/pytest_dist.log
for n in 4 8; do
for ch in 1 4; do
NIFTYB200_SLAB_CHUNKS=$ch timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench_n${n}_c$ch.json 2> gpurun_out/bench_n${n}_c$ch.err; echo "bench n=$n chunks=$ch rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_n${n}_c$ch.json')); print($n, $ch, d['value'], d['ms_per_step'], d['extra']['frac_of_combined_roofline'], d['clocks'])"
done
done

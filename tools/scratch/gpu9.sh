mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_distributed.py -q > gpurun_out/pytest_all.log 2>&1; echo "pytest rc=$?"; grep -E "^E  |passed|failed|skipped" gpurun_out/pytest_all.log | head -12
for n in 2 4 8; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "bench$n rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_n$n.json')); print($n, d['value'], d['ms_per_step'], d['extra'])"
done

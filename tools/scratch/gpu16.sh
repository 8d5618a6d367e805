mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_all.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu_all.log
timeout 600 python tools/prof_c5.py f32 2>&1 | tail -5 | tee gpurun_out/prof_c5.log
timeout 300 python bench.py --shape 512,512,512 --steps 20 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_c3.json 2>gpurun_out/bench_c3.err; echo "c3 rc=$?"; cut -c 1-300 gpurun_out/bench_c3.json
timeout 300 python bench.py --shape 4096,4096 --steps 50 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_c2.json 2>gpurun_out/bench_c2.err; echo "c2 rc=$?"; cut -c 1-300 gpurun_out/bench_c2.json
timeout 300 python bench.py --dtype f64 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_f64.json 2>gpurun_out/bench_f64.err; echo "f64 rc=$?"; cut -c 1-300 gpurun_out/bench_f64.json

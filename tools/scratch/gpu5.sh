mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -E "^E  .*Assertion|passed|failed" gpurun_out/pytest_gpu.log | head -8
python tools/prof_passes.py f32 1024 6 2>&1 | tee gpurun_out/passes_f32.log
python tools/prof_passes.py f64 1024 4 2>&1 | tee gpurun_out/passes_f64.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_f32.json 2> gpurun_out/bench_f32.err; echo "bench rc=$?"; cat gpurun_out/bench_f32.json

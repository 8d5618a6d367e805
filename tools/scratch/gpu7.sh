mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -E "^E  .*Assertion|passed|failed" gpurun_out/pytest_gpu.log | head -8
python tools/prof_passes.py f32 1024 6 2>&1 | tee gpurun_out/passes_f32.log
timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_f32.json 2> gpurun_out/bench_f32.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/bench_f32.json')); print(d['value'], d['ms_per_step'], d['extra']['cg_hbm_frac_of_measured']); print(json.dumps(d['extra']['kernel_ms'],indent=0)); print(d['extra']['fwd_applies_per_s'], d['extra']['adj_applies_per_s'], d['e2e'])"

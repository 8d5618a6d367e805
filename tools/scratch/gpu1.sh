mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/gpu.txt; nproc >> gpurun_out/gpu.txt; free -g >> gpurun_out/gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_f32.json 2> gpurun_out/bench_f32.err; echo "bench rc=$?"; cat gpurun_out/bench_f32.json; tail -5 gpurun_out/bench_f32.err
timeout 600 python bench.py --steps 5 --warmup 3 --dtype f64 --no-cpu-baseline --no-e2e > gpurun_out/bench_f64.json 2> gpurun_out/bench_f64.err; echo "bench64 rc=$?"; cat gpurun_out/bench_f64.json
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu1.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:StridedPass -s 6 -c 2 -o gpurun_out/prof_strided python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"

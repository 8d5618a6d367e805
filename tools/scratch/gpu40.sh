mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest40.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest40.log | cut -c1-200
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench40.json 2> gpurun_out/bench40.err; echo "bench rc=$?"; cut -c1-250 gpurun_out/bench40.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --steps 10 --warmup 3 --dtype f64 --no-cpu-baseline > gpurun_out/bench40_f64.json 2>/dev/null; echo "f64 rc=$?"; cut -c1-200 gpurun_out/bench40_f64.json
timeout 300 python bench.py --steps 20 --warmup 3 --shape 512,512,512 --no-cpu-baseline > gpurun_out/bench40_c3.json 2>/dev/null; echo "c3 rc=$?"; cut -c1-200 gpurun_out/bench40_c3.json

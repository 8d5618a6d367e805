mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest38.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest38.log | cut -c1-200
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench38.json 2> gpurun_out/bench38.err; echo "bench rc=$?"; cut -c1-250 gpurun_out/bench38.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches38.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu38.log 2>&1; echo "ncu rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --dtype f64 --no-cpu-baseline > gpurun_out/bench38_f64.json 2>/dev/null; echo "f64 rc=$?"; cut -c1-200 gpurun_out/bench38_f64.json
timeout 300 python bench.py --steps 20 --warmup 3 --shape 512,512,512 --no-cpu-baseline > gpurun_out/bench38_c3.json 2>/dev/null; echo "c3 rc=$?"; cut -c1-200 gpurun_out/bench38_c3.json
timeout 300 python bench.py --steps 50 --warmup 5 --shape 4096,4096 --no-cpu-baseline > gpurun_out/bench38_c2.json 2>/dev/null; echo "c2 rc=$?"; cut -c1-200 gpurun_out/bench38_c2.json

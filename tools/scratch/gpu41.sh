mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest41.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest41.log | cut -c1-200
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench41.json 2> gpurun_out/bench41.err; echo "bench rc=$?"; cut -c1-250 gpurun_out/bench41.json
NIFTYB200_FUSED_AXPBY=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench41_unfused.json 2>/dev/null; echo "unfused rc=$?"; cut -c1-250 gpurun_out/bench41_unfused.json
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1

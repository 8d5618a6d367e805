mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest44.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest44.log | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | cut -c1-200

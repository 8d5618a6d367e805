mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest20.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest20.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench20.json 2> gpurun_out/bench20.err; echo "bench rc=$?"; cat gpurun_out/bench20.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench20_ref.json 2> gpurun_out/bench20_ref.err; echo "ref rc=$?"; cat gpurun_out/bench20_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches20.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu20.log 2>&1; echo "ncu rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2

mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench42.json 2> gpurun_out/bench42.err; echo "bench rc=$?"; cut -c1-250 gpurun_out/bench42.json; tail -3 gpurun_out/bench42.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches42.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu42.log 2>&1; echo "ncu rc=$?"

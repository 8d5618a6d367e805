mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -E "AssertionError|passed|failed" gpurun_out/pytest_gpu.log | head -8
echo "== default"; python tools/prof_passes.py f32 1024 3 2>&1 | tee gpurun_out/passes_default.log
echo "== minb3"; NIFTYB200_LIB=$PWD/joss-nifty_b200/niftyb200/libniftyb200_minb3.so python tools/prof_passes.py f32 1024 3 2>&1 | tee gpurun_out/passes_minb3.log
echo "== plan a (16,4,16)"; NB2_PLAN10=a python tools/prof_passes.py f32 1024 3 2>&1 | head -4
echo "== plan b (8,16,8)"; NB2_PLAN10=b python tools/prof_passes.py f32 1024 3 2>&1 | head -4
echo "== logw 4"; NB2_STRIDED_LOGW=4 python tools/prof_passes.py f32 1024 3 2>&1 | head -4
echo "== logw 4 plan a"; NB2_STRIDED_LOGW=4 NB2_PLAN10=a python tools/prof_passes.py f32 1024 3 2>&1 | head -4
echo "== logw 2"; NB2_STRIDED_LOGW=2 python tools/prof_passes.py f32 1024 3 2>&1 | head -4
echo "== f64 default"; python tools/prof_passes.py f64 1024 2 2>&1 | tee gpurun_out/passes_f64.log
python tools/prof_passes.py f32 1024 0 > gpurun_out/plain2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:nb2_kernel -s 14 -c 10 -o gpurun_out/prof_r1a python tools/prof_passes.py f32 1024 0 > gpurun_out/ncu3.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/*.ncu-rep

mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -k "not full_size and not c5" > gpurun_out/pytest18.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest18.log
run() { echo "== $*" | tee -a gpurun_out/passes18.log; env "$@" timeout 300 python tools/prof_passes.py f32 1024 5 2>&1 | tail -10 | tee -a gpurun_out/passes18.log; }
run A=1
run NB2_PLAN10=a
run NB2_PLAN10=b
echo "== 512" | tee -a gpurun_out/passes18.log; timeout 300 python tools/prof_passes.py f32 512 5 2>&1 | tail -10 | tee -a gpurun_out/passes18.log
echo "== f64" | tee -a gpurun_out/passes18.log; timeout 300 python tools/prof_passes.py f64 1024 5 2>&1 | tail -10 | tee -a gpurun_out/passes18.log

mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -k "not full_size" > gpurun_out/pytest_pipe.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_pipe.log
run() { echo "== $*" | tee -a gpurun_out/passes_pipe.log; env "$@" timeout 300 python tools/prof_passes.py f32 1024 5 2>&1 | tail -12 | tee -a gpurun_out/passes_pipe.log; }
run NB2_PIPE=0
run NB2_PIPE=1
run NB2_STAGES=2
run NB2_ROWS_G=4
run NB2_ROWS_G=4 NB2_THREADS=256
run NB2_ROWS_G=8 NB2_THREADS=512 NB2_STAGES=2
run NB2_STRIDED_LOGW=3 NB2_ROWS_G=2 NB2_THREADS=128

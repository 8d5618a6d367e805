mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -k "not full_size and not c5" > gpurun_out/pytest17.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest17.log
export PROF_ONLY=rows,rows+mul,rows+field+add+dot
for n in 1024 512; do echo "== n=$n" | tee -a gpurun_out/passes17.log; timeout 300 python tools/prof_passes.py f32 $n 5 2>&1 | tail -3 | tee -a gpurun_out/passes17.log; done
for g in 2 8; do echo "== n=1024 G=$g" | tee -a gpurun_out/passes17.log; NB2_ROWS_G=$g timeout 300 python tools/prof_passes.py f32 1024 5 2>&1 | tail -3 | tee -a gpurun_out/passes17.log; done
echo "== f64 1024" | tee -a gpurun_out/passes17.log; timeout 300 python tools/prof_passes.py f64 1024 5 2>&1 | tail -3 | tee -a gpurun_out/passes17.log

mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "tma or metric or cf_ or cg_parity or full_size" > gpurun_out/pytest39.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest39.log | cut -c1-200
export PROF_ONLY=strided0,strided0+field
for v in 1 0; do echo "== NB2_TMA_FX=$v"; NB2_TMA_FX=$v timeout 300 python tools/prof_passes.py f32 1024 5 2>&1 | tail -2; done | tee gpurun_out/passes39.log
for v in 1 0; do echo "== f64 NB2_TMA_FX=$v"; NB2_TMA_FX=$v timeout 300 python tools/prof_passes.py f64 1024 5 2>&1 | tail -2; done | tee -a gpurun_out/passes39.log
for v in 1 0; do echo "== 512 NB2_TMA_FX=$v"; NB2_TMA_FX=$v timeout 300 python tools/prof_passes.py f32 512 10 2>&1 | tail -2; done | tee -a gpurun_out/passes39.log

mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "tma_strided" > gpurun_out/pytest27.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest27.log | cut -c1-250
export PROF_ONLY=strided0,strided1,strided0+field,strided0+amp
for v in 1 0; do echo "== NB2_TMA=$v"; NB2_TMA=$v timeout 300 python tools/prof_passes.py f32 1024 5 2>&1 | tail -4; done | tee gpurun_out/passes27.log
for v in 1 0; do echo "== 512 NB2_TMA=$v"; NB2_TMA=$v timeout 300 python tools/prof_passes.py f32 512 10 2>&1 | tail -4; done | tee -a gpurun_out/passes27.log
for v in 1 0; do echo "== f64 NB2_TMA=$v"; NB2_TMA=$v timeout 300 python tools/prof_passes.py f64 1024 5 2>&1 | tail -4; done | tee -a gpurun_out/passes27.log

export PROF_ONLY=rows,rows+mul,rows+field+add+dot
for v in 1 0; do echo "== NB2_ROWS_SPEC=$v"; NB2_ROWS_SPEC=$v timeout 200 python tools/prof_passes.py f32 1024 5 2>&1 | tail -3; done | tee gpurun_out/passes45.log

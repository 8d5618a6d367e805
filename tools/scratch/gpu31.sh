mkdir -p gpurun_out
export PROF_ONLY=rows,rows+mul,rows+field+add+dot
run() { echo "== $*"; env "$@" timeout 300 python tools/prof_passes.py f32 1024 5 2>&1 | tail -3; }
{ run NB2_ROWS_LA=0; run A=1; run NB2_ROWS_LA=370; run NB2_ROWS_LA=1480; run NB2_ROWS_LA=2960; } | tee gpurun_out/passes31.log

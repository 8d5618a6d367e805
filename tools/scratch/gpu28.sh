mkdir -p gpurun_out
export PROF_ONLY=strided0,strided1,strided0+field
run() { echo "== $*"; env "$@" timeout 300 python tools/prof_passes.py f32 1024 5 2>&1 | tail -3; }
{ run A=1; run NB2_TMA_L2P=2; run NB2_TMA_L2P=3; run NB2_TMA_HINT=1; run NB2_TMA_HINT=1 NB2_TMA_L2P=3; } | tee gpurun_out/passes28.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench28.json 2> gpurun_out/bench28.err; echo "bench rc=$?"; cut -c1-1800 gpurun_out/bench28.json

mkdir -p gpurun_out
export PROF_ONLY=rows,rows+mul,rows+field+add+dot
{ echo "== remap (default)"; timeout 300 python tools/prof_passes.py f32 1024 5 2>&1 | tail -3
  echo "== no remap"; NIFTYB200_LIB=$PWD/joss-nifty_b200/niftyb200/libniftyb200_noremap.so timeout 300 python tools/prof_passes.py f32 1024 5 2>&1 | tail -3
  echo "== 512 remap"; timeout 300 python tools/prof_passes.py f32 512 10 2>&1 | tail -3
  echo "== 512 no remap"; NIFTYB200_LIB=$PWD/joss-nifty_b200/niftyb200/libniftyb200_noremap.so timeout 300 python tools/prof_passes.py f32 512 10 2>&1 | tail -3; } | tee gpurun_out/passes37.log

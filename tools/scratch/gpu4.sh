mkdir -p gpurun_out
L=$PWD/joss-nifty_b200/niftyb200
run() { echo "== $1"; shift; env "$@" python tools/prof_passes.py f32 1024 6 2>&1 | grep -v Warn; }
export PROF_ONLY=strided0,strided1,rows,rows+mul
run "base (minb2)" A=1
run "twpow" NIFTYB200_LIB=$L/libniftyb200_twpow.so
run "minb3" NIFTYB200_LIB=$L/libniftyb200_minb3.so
run "twpow3" NIFTYB200_LIB=$L/libniftyb200_twpow3.so
run "twpow plan10a plan9a" NIFTYB200_LIB=$L/libniftyb200_twpow.so NB2_PLAN10=a NB2_PLAN9=a
run "twpow plan10b plan9b" NIFTYB200_LIB=$L/libniftyb200_twpow.so NB2_PLAN10=b NB2_PLAN9=b
run "twpow plan9c G4" NIFTYB200_LIB=$L/libniftyb200_twpow.so NB2_PLAN9=c NB2_ROWS_G=4
run "twpow logw4" NIFTYB200_LIB=$L/libniftyb200_twpow.so NB2_STRIDED_LOGW=4
run "twpow logw4 plan10a" NIFTYB200_LIB=$L/libniftyb200_twpow.so NB2_STRIDED_LOGW=4 NB2_PLAN10=a
run "twpow G4" NIFTYB200_LIB=$L/libniftyb200_twpow.so NB2_ROWS_G=4
run "twpow G2 plan9a" NIFTYB200_LIB=$L/libniftyb200_twpow.so NB2_ROWS_G=2 NB2_PLAN9=a
run "twpow G4 plan9a" NIFTYB200_LIB=$L/libniftyb200_twpow.so NB2_ROWS_G=4 NB2_PLAN9=a
run "twpow G4 plan9b" NIFTYB200_LIB=$L/libniftyb200_twpow.so NB2_ROWS_G=4 NB2_PLAN9=b

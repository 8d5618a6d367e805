mkdir -p gpurun_out
export PROF_ONLY=strided0,strided1,rows,rows+mul,strided0+field,rows+field+add+dot
run() { echo "== $*" | tee -a gpurun_out/passes19.log; env "$@" timeout 300 python tools/prof_passes.py ${DT:-f32} ${NN:-1024} 5 2>&1 | tail -6 | tee -a gpurun_out/passes19.log; }
run NB2_TWTABLE=1
run NB2_TWTABLE=0
NN=512 run NB2_TWTABLE=1
NN=512 run NB2_TWTABLE=0
DT=f64 run NB2_TWTABLE=1
DT=f64 run NB2_TWTABLE=0

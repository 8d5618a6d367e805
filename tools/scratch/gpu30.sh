mkdir -p gpurun_out
export PROF_ONLY=strided0,strided1,strided0+field,rows+mul,rows+field+add+dot
PROF_NCU=1 timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:nb2_kernel -c 5 -f -o gpurun_out/r1_full_v3 python tools/prof_passes.py f32 1024 0 > gpurun_out/ncu30.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu30.log; ls -la gpurun_out/*.ncu-rep

mkdir -p gpurun_out
NB2_L2_STREAMS=1 timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:nb2_kernel --launch-skip 40 -c 8 --csv --log-file gpurun_out/ncu26_l2.csv python tools/scratch/dbg_l2.py 1024 8 > gpurun_out/ncu26.log 2>&1; echo rc=$?
grep '^"' gpurun_out/ncu26_l2.csv | cut -d, -f5,9,13- | cut -c1-230 | head -40

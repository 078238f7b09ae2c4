bash tools/gpu_check_r02.sh r02z > gpurun_out/r02z_check.log 2>&1; tail -12 gpurun_out/r02z_check.log | cut -c1-300
M=gpu__time_duration.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r02z_activity_c5chunk.csv python tools/ncu_target.py --shape 4x512x512 --frames 3 > /dev/null 2>&1
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r02z_activity_1080p.csv python tools/ncu_target.py --shape 1x270x480 --frames 3 --compose > /dev/null 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv3x3_pair_kernel -s 30 -c 4 -o gpurun_out/r02z_full python tools/ncu_target.py --shape 4x512x512 --frames 2 > gpurun_out/r02z_ncu_full.log 2>&1
tail -2 gpurun_out/r02z_ncu_full.log; ls -la gpurun_out/r02z_full.ncu-rep

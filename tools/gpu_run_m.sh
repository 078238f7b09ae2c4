mkdir -p gpurun_out
M=gpu__time_duration.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum
timeout 300 python tools/ncu_target.py --shape 4x512x512 --frames 3 > /dev/null 2>&1 && timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r02m_activity_c5chunk.csv python tools/ncu_target.py --shape 4x512x512 --frames 3 > /dev/null 2>&1
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r02m_activity_1080p.csv python tools/ncu_target.py --shape 1x270x480 --frames 3 --compose > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none -k regex:conv3x3_pair_kernel -s 29 -c 3 -o gpurun_out/r02m_trunk python tools/ncu_target.py --shape 4x512x512 --frames 2 > gpurun_out/r02m_ncu_trunk.log 2>&1
tail -2 gpurun_out/r02m_ncu_trunk.log

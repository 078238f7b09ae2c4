mkdir -p gpurun_out
M=gpu__time_duration.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum
timeout 300 python tools/ncu_target.py --shape 4x512x512 --frames 3 > /dev/null 2>&1 && timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r02k_activity_c5chunk.csv python tools/ncu_target.py --shape 4x512x512 --frames 3 > /dev/null 2>&1
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r02k_activity_1080p.csv python tools/ncu_target.py --shape 1x270x480 --frames 3 --compose > /dev/null 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,dram__bytes_read.sum,dram__bytes_write.sum,l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum --clock-control none -k regex:pack_recurrent -c 3 --csv --log-file gpurun_out/r02k_pack_smooth.csv python tools/ncu_target.py --shape 4x512x512 --frames 3 --smooth-flow > /dev/null 2>&1
grep -E "pack_recurrent" gpurun_out/r02k_pack_smooth.csv | awk -F'","' '{print $(NF-2), $(NF)}' | tail -6
timeout 120 python __graft_entry__.py smoke

mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv3x3_final2 -s 1 -c 1 -o gpurun_out/r02f_final2 python tools/ncu_target.py --shape 4x512x512 --frames 2 > gpurun_out/r02f_ncu.log 2>&1
tail -2 gpurun_out/r02f_ncu.log

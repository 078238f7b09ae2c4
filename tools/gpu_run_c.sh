mkdir -p gpurun_out
for v in 1 2; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pack_recurrent -c 2 -o gpurun_out/r02c_pack_v$v python tools/ncu_target.py --shape 4x512x512 --frames 2 --tunable pack_variant=$v > gpurun_out/r02c_ncu_pack_v$v.log 2>&1
done
timeout 600 ncu --set full --clock-control none -k regex:conv3x3_pair_kernel -s 29 -c 3 -o gpurun_out/r02c_trunk python tools/ncu_target.py --shape 4x512x512 --frames 2 > gpurun_out/r02c_ncu_trunk.log 2>&1
tail -3 gpurun_out/r02c_ncu_pack_v2.log gpurun_out/r02c_ncu_trunk.log

mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02i_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r02i_pytest.log
for v in 2 3; do
timeout 600 ncu --metrics gpu__time_duration.sum,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed --clock-control none -k regex:pack_recurrent -c 3 --csv --log-file gpurun_out/r02i_pack_v$v.csv python tools/ncu_target.py --shape 4x512x512 --frames 3 --tunable pack_variant=$v > /dev/null 2>&1
grep -E "gpu__time|wavefronts" gpurun_out/r02i_pack_v$v.csv | awk -F'","' '{print $(NF-2), $(NF)}' | tail -4
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:pack_recurrent -c 3 --csv --log-file gpurun_out/r02i_pack_1080p.csv python tools/ncu_target.py --shape 1x270x480 --frames 3 > /dev/null 2>&1
grep -E "gpu__time" gpurun_out/r02i_pack_1080p.csv | awk -F'","' '{print $(NF-2), $(NF)}' | tail -2

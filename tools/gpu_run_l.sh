mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02l_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/r02l_pytest.log
timeout 300 python tools/pair_experiments.py 2>&1 | grep "dbg=0 scale=1"
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02l_bench.json 2> gpurun_out/r02l_bench.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/r02l_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['latency_1080p']['p50_ms'], d['configs1']['value'], d['configs3']['value'], d['clocks'], d['roofline']['whole_conv_stack']['frac_of_burst'])"
M=gpu__time_duration.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r02m_activity_c5chunk.csv python tools/ncu_target.py --shape 4x512x512 --frames 3 > /dev/null 2>&1
timeout 900 ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r02m_activity_1080p.csv python tools/ncu_target.py --shape 1x270x480 --frames 3 --compose > /dev/null 2>&1

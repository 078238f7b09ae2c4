mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02g_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/r02g_pytest.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02g_bench.json 2> gpurun_out/r02g_bench.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/r02g_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['latency_1080p']['p50_ms'], d['configs1']['value'], d['configs3']['value'], d['clocks'], d['roofline']['frac'], d['roofline']['whole_conv_stack'])"
timeout 300 python tools/ncu_target.py --shape 4x512x512 --frames 3 > /dev/null 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02g_launches_c5chunk.csv python tools/ncu_target.py --shape 4x512x512 --frames 3 > /dev/null 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02g_launches_1080p.csv python tools/ncu_target.py --shape 1x270x480 --frames 3 --compose > /dev/null 2>&1

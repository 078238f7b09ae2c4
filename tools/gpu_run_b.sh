mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02b_pytest.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/r02b_pytest.log
timeout 900 python tools/parity_margin.py --shape 2x256x256 > gpurun_out/r02b_parity_margin.json 2> gpurun_out/r02b_parity_margin.txt; echo "margin exit $?"; tail -12 gpurun_out/r02b_parity_margin.txt
timeout 300 python tools/ncu_target.py --shape 4x512x512 --frames 3 > /dev/null 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02b_launches_c5chunk.csv python tools/ncu_target.py --shape 4x512x512 --frames 3 > /dev/null 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02b_launches_1080p.csv python tools/ncu_target.py --shape 1x270x480 --frames 3 --compose > /dev/null 2>&1
grep -E "pack_recurrent|compose" gpurun_out/r02b_launches_c5chunk.csv gpurun_out/r02b_launches_1080p.csv | awk -F'","' '{print $5, $(NF)}' | tail -8
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02b_bench.json 2> gpurun_out/r02b_bench.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/r02b_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['latency_1080p']['p50_ms'], d['clocks'])"

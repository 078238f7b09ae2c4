mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r02l_pytest.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/r02l_pytest.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r02l_bench.json 2> gpurun_out/r02l_bench.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/r02l_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['latency_1080p']['p50_ms'], d['configs1']['value'], d['configs3']['value'], d['clocks'], d['roofline']['whole_conv_stack']['frac_of_burst'])"
timeout 900 python tools/recurrence_drift.py > gpurun_out/r02l_recurrence_drift.json 2> gpurun_out/r02l_recurrence_drift.txt; grep -E "hybrid|torch_cuda" gpurun_out/r02l_recurrence_drift.txt
timeout 120 python __graft_entry__.py smoke

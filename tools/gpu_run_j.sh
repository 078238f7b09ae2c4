mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02j_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r02j_pytest.log
timeout 600 python tools/pdl_experiments.py 2>&1 | grep "pdl_small=" | head -4
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02j_bench.json 2> gpurun_out/r02j_bench.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/r02j_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['latency_1080p']['p50_ms'], d['latency_1080p']['p99_ms'], d['latency_1080p']['from_render']['p50_ms'], d['clocks'])"

#!/bin/bash
# One gpurun call: GPU parity tests, bench, ncu launch list, ncu full capture of the dominant conv kernel.
# usage: tools/gpu_check.sh <tag>
TAG=${1:-r01}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 120 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; tail -1 gpurun_out/${TAG}_smoke.log
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err
if [ "$2" != "noncu" ]; then
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/${TAG}_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv3x3_ -s 58 -c 2 \
    -o gpurun_out/${TAG}_conv_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-latency > gpurun_out/${TAG}_ncu2.log 2>&1
fi
tail -3 gpurun_out/${TAG}_pytest.log; cat gpurun_out/${TAG}_bench.json; cat gpurun_out/${TAG}_bench_ref.json

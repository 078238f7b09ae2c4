#!/bin/bash
# One gpurun call: smoke, GPU parity tests, bench (both arms), optional extras.
# usage: tools/gpu_check_r02.sh <tag> [drift] [ncu]
TAG=${1:-r02}
shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
nvidia-smi topo -m >> gpurun_out/${TAG}_smi.txt 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; tail -1 gpurun_out/${TAG}_smoke.log
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
tail -40 gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"; tail -5 gpurun_out/${TAG}_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err
for what in "$@"; do
case $what in
drift)
  timeout 900 python tools/recurrence_drift.py > gpurun_out/${TAG}_recurrence_drift.json 2> gpurun_out/${TAG}_recurrence_drift.txt
  echo "drift exit $?"; tail -45 gpurun_out/${TAG}_recurrence_drift.txt ;;
ncu)
  timeout 300 python tools/ncu_target.py --shape 4x512x512 --frames 3 > gpurun_out/${TAG}_target.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/${TAG}_launches_c5chunk.csv python tools/ncu_target.py --shape 4x512x512 --frames 3 > gpurun_out/${TAG}_ncu1.log 2>&1
  timeout 300 python tools/ncu_target.py --shape 1x270x480 --frames 3 --compose > gpurun_out/${TAG}_target2.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/${TAG}_launches_1080p.csv python tools/ncu_target.py --shape 1x270x480 --frames 3 --compose > gpurun_out/${TAG}_ncu1b.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"conv3x3_|pack_recurrent" -s 80 -c 40 \
      -o gpurun_out/${TAG}_full python tools/ncu_target.py --shape 4x512x512 --frames 3 > gpurun_out/${TAG}_ncu2.log 2>&1
  ;;
esac
done
cat gpurun_out/${TAG}_bench.json; cat gpurun_out/${TAG}_bench_ref.json

mkdir -p gpurun_out
N=${1:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r02_bench_n$N.out 2> gpurun_out/r02_bench_n$N.err; echo "bench n$N exit $?"; tail -2 gpurun_out/r02_bench_n$N.err
grep '^{' gpurun_out/r02_bench_n$N.out | tail -1 > gpurun_out/r02_bench_n$N.json
python -c "
import json,sys; d=json.load(open('gpurun_out/r02_bench_n$N.json')); e=d['e2e']; print('value',d['value'],'ms',d['ms_per_step'],'e2e',e['value'],'ceil',e['d2h_ceiling_gbs'],'frac',e['d2h_frac_of_ceiling'],'fp16',e['fp16_out']['value'],'disp',e['display']['value'],e['affinity'], d['clocks'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/r02_bench_ref_n$N.json 2>> gpurun_out/r02_bench_n$N.err; echo "ref exit $?"
nvidia-smi topo -m > gpurun_out/r02_topo_n$N.txt 2>&1; lscpu | head -20 >> gpurun_out/r02_topo_n$N.txt

mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02d_bench_n2.json 2> gpurun_out/r02d_bench_n2.err; echo "bench n2 exit $?"; tail -3 gpurun_out/r02d_bench_n2.err; cat gpurun_out/r02d_bench_n2.json | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['d2h_ceiling_gbs'], d['e2e']['d2h_frac_of_ceiling'], d['e2e']['fp16_out']['value'], d['e2e']['display']['value'], d['e2e']['affinity'])"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r02d_bench_ref_n2.json 2>> gpurun_out/r02d_bench_n2.err; echo "ref n2 exit $?"; wc -l gpurun_out/r02d_bench_ref_n2.json
python -m pytest tests -m gpu -q -k "two_devices" 2>&1 | tail -3
nvidia-smi topo -m > gpurun_out/r02d_topo.txt 2>&1

mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2_v15_bench_8gpu.json 2> gpurun_out/r2_v15_bench_8gpu.err; echo "rc $?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_v15_bench_8gpu.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['n_gpus'], d['per_launch_ms'])
print(json.dumps(d.get('secondary'))[:1500])
print(d.get('mc_stats'))
"
tail -5 gpurun_out/r2_v15_bench_8gpu.err

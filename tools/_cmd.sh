mkdir -p gpurun_out
brief() { python -c "import json,sys; d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], '%.4g' % d['value'], d['per_launch_ms'])" $1 $2; }
for g in 1 64 128; do VIEKF_GROUP=$g python bench.py --filters 65536 --features 8 --no-cpu --no-secondary > gpurun_out/n8_g$g.json 2>/dev/null; brief gpurun_out/n8_g$g.json N8_group$g; done
for n in 1 2 4 6; do python bench.py --filters 65536 --features $n --no-cpu --no-secondary > gpurun_out/n$n.json 2>/dev/null; brief gpurun_out/n$n.json N$n; done

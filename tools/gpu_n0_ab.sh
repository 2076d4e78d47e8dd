#!/bin/bash
# A/B of build_ab/lib_*.so builds of the N = 0 executor at the configs[1] size; optional ncu --set full capture of the in-tree build.
TAG=${1:-r2_oct_ab}
mkdir -p gpurun_out
brief() { python -c "import json,sys; d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], '%.4g' % d['value'], d['per_launch_ms'])" $1 $2; }
python bench.py --filters 65536 --features 0 --no-cpu --no-secondary > gpurun_out/${TAG}_base.json 2>/dev/null; brief gpurun_out/${TAG}_base.json in-tree | tee gpurun_out/${TAG}.txt
for so in build_ab/lib_*.so; do
  VIEKF_LIB=$PWD/$so python bench.py --filters 65536 --features 0 --no-cpu --no-secondary > gpurun_out/${TAG}_tmp.json 2>/dev/null; brief gpurun_out/${TAG}_tmp.json $so | tee -a gpurun_out/${TAG}.txt
done
if [ -n "$2" ]; then
  ncu --set full --clock-control none --import-source on -k regex:viekf_octet_kernel -s 3 -c 1 -f -o gpurun_out/${TAG}_octet python bench.py --filters 65536 --features 0 --steps 1 --warmup 3 --no-cpu --no-secondary > gpurun_out/${TAG}_ncu_full.log 2>&1; echo "ncu full rc $?"
fi

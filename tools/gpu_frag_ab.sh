#!/bin/bash
# A/B of build_ab/lib_*.so builds of the fragment executor at 65,536 x N (default N = 8) against the in-tree build.
TAG=${1:-r2_frag_ab}; NF=${2:-8}
mkdir -p gpurun_out
brief() { python -c "import json,sys; d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], '%.4g' % d['value'], d['per_launch_ms'])" $1 $2; }
python bench.py --filters 65536 --features $NF --no-cpu --no-secondary > gpurun_out/${TAG}_base.json 2>/dev/null; brief gpurun_out/${TAG}_base.json in-tree | tee gpurun_out/${TAG}.txt
for so in build_ab/lib_*.so; do
  VIEKF_LIB=$PWD/$so python bench.py --filters 65536 --features $NF --no-cpu --no-secondary > gpurun_out/${TAG}_tmp.json 2>/dev/null; brief gpurun_out/${TAG}_tmp.json $so | tee -a gpurun_out/${TAG}.txt
done
python bench.py --filters 65536 --features $NF --no-cpu --no-secondary > gpurun_out/${TAG}_base2.json 2>/dev/null; brief gpurun_out/${TAG}_base2.json in-tree-again | tee -a gpurun_out/${TAG}.txt

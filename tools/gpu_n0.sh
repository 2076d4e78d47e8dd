#!/bin/bash
# One gpurun call for the N = 0 executor: parity tests that touch N = 0, then the configs[1]-size bench under both mappings.
TAG=${1:-r2_oct}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "N0 or all_event_kinds or config_variants or baseline_sizes or queue or no_alloc or per_call or edge or statistics" > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc $?"
tail -15 gpurun_out/${TAG}_pytest.log
python bench.py --filters 65536 --features 0 --no-cpu --no-secondary > gpurun_out/${TAG}_bench_n0.json 2> gpurun_out/${TAG}_bench_n0.err; echo "bench rc $?"
VIEKF_GROUP=32 python bench.py --filters 65536 --features 0 --no-cpu --no-secondary > gpurun_out/${TAG}_bench_n0_g32.json 2>/dev/null
for f in gpurun_out/${TAG}_bench_n0.json gpurun_out/${TAG}_bench_n0_g32.json; do python -c "import json,sys; d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], d[\"value\"], d[\"per_launch_ms\"], d[\"roofline\"][\"kernel\"])" $f; done

#!/bin/bash
# One gpurun call: GPU tests, smoke, both bench arms, ncu launch list, DRAM traffic of the tile kernel at the bench size and one
# --set full capture of it.
TAG=${1:-r2_v12}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc $?" | tee -a gpurun_out/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc $?"
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc $?"
python bench.py --impl reference > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "ref rc $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/${TAG}_ncu_list.log 2>&1; echo "ncu list rc $?"
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:viekf_tile_kernel -s 2 -c 1 --csv --log-file gpurun_out/${TAG}_traffic.csv python bench.py --steps 1 --warmup 2 --no-cpu --no-secondary > gpurun_out/${TAG}_ncu_traffic.log 2>&1; echo "ncu traffic rc $?"
ncu --set full --clock-control none --import-source on -k regex:viekf_tile_kernel -s 1 -c 1 -f -o gpurun_out/${TAG}_tile python bench.py --filters 16384 --steps 1 --warmup 1 --no-cpu --no-secondary > gpurun_out/${TAG}_ncu_full.log 2>&1; echo "ncu full rc $?"
tail -3 gpurun_out/${TAG}_pytest.log; cat gpurun_out/${TAG}_bench.json | cut -c1-600

#!/bin/bash
# One gpurun call: GPU tests, smoke, both bench arms, ncu launch list, DRAM traffic of the tile kernel at the bench size and one
# --set full capture of each executor (tile, fragment, octet).
TAG=${1:-r2_v18}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc $?" | tee -a gpurun_out/${TAG}_pytest.log
python -m pytest tests/test_gpu_bench_tape.py -x -q -s -k bench_tape_parity 2>&1 | grep -E "worst|passed|failed" > gpurun_out/${TAG}_bench_tape_parity.txt; cat gpurun_out/${TAG}_bench_tape_parity.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc $?"
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc $?"
python bench.py --impl reference > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "ref rc $?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/${TAG}_ncu_list.log 2>&1; echo "ncu list rc $?"
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:viekf_tile_kernel -s 2 -c 1 --csv --log-file gpurun_out/${TAG}_traffic.csv python bench.py --steps 1 --warmup 2 --no-cpu --no-secondary > gpurun_out/${TAG}_ncu_traffic.log 2>&1; echo "ncu traffic rc $?"
ncu --set full --clock-control none --import-source on -k regex:viekf_tile_kernel -s 1 -c 1 -f -o gpurun_out/${TAG}_tile python bench.py --filters 16384 --steps 1 --warmup 1 --no-cpu --no-secondary > gpurun_out/${TAG}_ncu_full.log 2>&1; echo "ncu full rc $?"
ncu --set full --clock-control none --import-source on -k regex:viekf_frag_kernel -s 3 -c 1 -f -o gpurun_out/${TAG}_frag python bench.py --filters 65536 --features 8 --steps 1 --warmup 3 --no-cpu --no-secondary > gpurun_out/${TAG}_ncu_frag.log 2>&1; echo "ncu frag rc $?"
ncu --set full --clock-control none --import-source on -k regex:viekf_octet_kernel -s 3 -c 1 -f -o gpurun_out/${TAG}_octet python bench.py --filters 65536 --features 0 --steps 1 --warmup 3 --no-cpu --no-secondary > gpurun_out/${TAG}_ncu_octet.log 2>&1; echo "ncu octet rc $?"
tail -3 gpurun_out/${TAG}_pytest.log; cat gpurun_out/${TAG}_bench.json | cut -c1-400

mkdir -p gpurun_out
python bench.py --filters 65536 --features 0 --no-cpu --no-secondary > gpurun_out/r2_oct2_bench_n0.json 2>/dev/null; cut -c1-300 gpurun_out/r2_oct2_bench_n0.json
python bench.py --filters 262144 --features 0 --no-cpu --no-secondary > gpurun_out/r2_oct2_bench_n0_262144.json 2>/dev/null; cut -c1-300 gpurun_out/r2_oct2_bench_n0_262144.json
ncu --set full --clock-control none --import-source on -k regex:viekf_octet_kernel -s 3 -c 1 -f -o gpurun_out/r2_oct2_octet python bench.py --filters 65536 --features 0 --steps 1 --warmup 3 --no-cpu --no-secondary > gpurun_out/r2_oct2_ncu_full.log 2>&1; echo "ncu full rc $?"

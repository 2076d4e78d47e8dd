mkdir -p gpurun_out
VIEKF_GROUP=64 ncu --set full --clock-control none --import-source on -k regex:viekf_tape_kernel -s 3 -c 1 -f -o gpurun_out/r2_g64_n8 python bench.py --filters 65536 --features 8 --steps 1 --warmup 3 --no-cpu --no-secondary > gpurun_out/r2_g64_n8_ncu.log 2>&1; echo "rc $?"
ncu --set full --clock-control none --import-source on -k regex:viekf_tape_kernel -s 3 -c 1 -f -o gpurun_out/r2_g32_n1 python bench.py --filters 65536 --features 1 --steps 1 --warmup 3 --no-cpu --no-secondary > gpurun_out/r2_g32_n1_ncu.log 2>&1; echo "rc $?"

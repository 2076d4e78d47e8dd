mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_v14_pytest.log 2>&1; echo "pytest rc $?"; tail -3 gpurun_out/r2_v14_pytest.log
python bench.py --no-cpu > gpurun_out/r2_v14_bench.json 2>gpurun_out/r2_v14_bench.err; python -c "
import json
d=json.loads(open('gpurun_out/r2_v14_bench.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']['value'], d['per_launch_ms'])
for k,v in d['secondary'].items(): print(k, v['value'])
"
VIEKF_LIB=$PWD/vi-ekf_b200/csrc/libviekf_b200_pht.so python tools/phase_timing.py 32768 > gpurun_out/r2_v14_phase_timing.txt 2>&1; cat gpurun_out/r2_v14_phase_timing.txt

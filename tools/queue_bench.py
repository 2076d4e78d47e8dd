"""Throughput of the streaming (rewind / replay queue) mode on BASELINE configs[0]'s call pattern (tuning / evidence helper): the
recorded call log of tests/queue_calls.py -- imu_cb -> propagate_state, feat_cb -> add_measurement per feature + handle_measurements,
N = 12, 250 Hz IMU / 30 Hz camera -- replayed on the engine in queue mode for B filters and on the CPU oracle (one filter per thread).
Prints one JSON line: calls/s, filter-steps/s (IMU samples x filters / wall time) for both."""
import json, os, sys, time
from concurrent.futures import ThreadPoolExecutor
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import __graft_entry__ as g
pkg = g.load_package()
from oracle import oracle as oracle_mod
from tests import queue_calls
abi = pkg.abi
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
n_imu = int(sys.argv[2]) if len(sys.argv) > 2 else 300
NF = 12
cfg = abi.default_config()
calls = queue_calls.config0_calls(pkg, cfg, NF, n_imu=n_imu)
n_prop = int((calls["kind"] == 0).sum())
oracle_mod.build()
# engine
bk = pkg.BatchVIEKF(cfg, B=B, N=NF)
bk.enable_queue(250, 200)
bk.perturb_x0(7)
queue_calls.replay(calls[:50], bk, abi, engine=True)  # warm-up (allocations, first launches)
bk.reset()
bk.perturb_x0(7)
bk.sync()
t0 = time.perf_counter()
adds, gated, n_upd = queue_calls.replay(calls, bk, abi, engine=True)
bk.sync()
t_gpu = time.perf_counter() - t0
launches = bk.launch_count()
# CPU oracle: one OracleVIEKF per thread, the same call log (the ctypes calls release the GIL)
threads = os.cpu_count() or 1
def run_one(_):
    f = oracle_mod.OracleVIEKF(cfg, NF)
    queue_calls.replay(calls, f, abi, engine=False)
    return 1
t0 = time.perf_counter()
with ThreadPoolExecutor(threads) as ex:
    list(ex.map(run_one, range(threads)))
t_cpu = time.perf_counter() - t0
print(json.dumps({"workload": f"BASELINE configs[0] call pattern (vi_ekf_test.cpp:24-33,58), N = {NF}, {len(calls)} calls of which {n_prop} propagate_state, {n_upd} replayed updates",
                  "engine": {"filters": B, "wall_s": t_gpu, "calls_per_s": len(calls) / t_gpu, "filter_steps_per_s": n_prop * B / t_gpu, "kernel_launches": launches},
                  "cpu_oracle": {"filters": threads, "threads": threads, "wall_s": t_cpu, "filter_steps_per_s": n_prop * threads / t_cpu,
                                 "note": "Python-driven ctypes calls into the C++ oracle, one filter per thread"}}))

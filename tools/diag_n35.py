"""Diagnostic (tuning helper): per-instance state / covariance distance to the oracle on the bench tape, with the gate counters."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import __graft_entry__ as g
pkg = g.load_package()
from oracle import oracle as oracle_mod
from tests.common import cov_rel_err, state_rel_err
N, B, steps, sample = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
if len(sys.argv) > 5: os.environ["VIEKF_TILE_FPC"] = sys.argv[5]
sim = pkg.sim
cfg, ev, idp, info, slices = sim.bench_workload(N, steps // sim.BENCH_IMU_PER_STEP)
nb = info["n_buckets"]
oracle_mod.build()
ref = oracle_mod.run_batch(cfg, N, ev, idp, count=sample, nthreads=32, seed=sim.BENCH_SEED, bias_range=sim.BENCH_BIAS, perturb_seed=sim.BENCH_SEED + 1, n_buckets=nb)
bk = pkg.BatchVIEKF(cfg, B=B, N=N)
bk.perturb_x0(sim.BENCH_SEED + 1)
res = bk.run_tape(ev, idp, results_first=0, results_count=sample, seed=sim.BENCH_SEED, bias_range=sim.BENCH_BIAS)
x, P = bk.get_state(0, sample), bk.get_covariance(0, sample)
ln, ids = bk.get_features(0, sample)
rows = []
for b in range(sample):
    nx, n = 17 + 5 * int(ln[b]), 16 + 3 * int(ln[b])
    ex = state_rel_err(x[b][:nx], ref["x"][b][:nx]); eP = cov_rel_err(P[b][:n, :n], ref["P"][b][:n, :n])
    Pr = ref["P"][b][:n, :n]; d = np.sqrt(np.diag(Pr)); C = Pr / np.outer(d, d)
    w = np.linalg.eigvalsh((C + C.T) / 2)
    rows.append((eP, ex, b, int(ref["flags"][b, 1]), float(ref["margins"][b, 0]), bool(np.array_equal(res[b], ref["results"][b])), w[-1] / max(w[0], 1e-300), float(np.abs(ref["x"][b][:3]).max())))
rows.sort(reverse=True)
print("eP ex inst gated margin results_equal cond(corr P) max|pos|")
for r in rows[:12]: print("%.2e %.2e %3d %5d %.1e %s %.2e %.2f" % r)
print("median eP %.2e" % np.median([r[0] for r in rows]))

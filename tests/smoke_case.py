"""One small invocation of the hot path on cuda:0 checked against the CPU oracle (used by __graft_entry__.smoke())."""
import numpy as np

from tests import tapes
from tests.common import REL_TOL


def run(pkg, verbose=False):
    """NF = 12 runs on the register-tile executor (the kernel family bench.py measures), NF = 6 on the fragment executor, NF = 0 on
    the lane-octet executor."""
    worst = (0.0, 0.0)
    for NF in (12, 6, 0):
        w = run_one(pkg, NF, verbose)
        worst = (max(worst[0], w[0]), max(worst[1], w[1]))
    return worst


def run_one(pkg, NF, verbose=False):
    from oracle import oracle
    oracle.build()
    abi = pkg.abi
    cfg = abi.default_config()
    B = 16
    ev, idp, f, nb = tapes.consistent_tape(oracle, cfg, NF, steps=36, seed=2)
    ref = oracle.run_batch(cfg, NF, ev, idp, count=B, nthreads=4, n_buckets=nb)
    bk = pkg.BatchVIEKF(cfg, B=B, N=NF, device=0)
    bk.stats_reset(nb)
    res = bk.run_tape(ev, idp, results_first=0, results_count=B)
    assert np.array_equal(res, ref["results"]), "smoke: per-event meas_result_t differ from the oracle"
    x, P = bk.get_state(), bk.get_covariance()
    ln, ids = bk.get_features()
    worst = (0.0, 0.0)
    for b in range(B):
        assert list(ids[b][: ln[b]]) == list(ref["ids"][b][: ln[b]])
        ex, eP = tapes.live_compare(x[b], P[b], ln[b], ref["x"][b], ref["P"][b], ref["len"][b], REL_TOL, f"smoke inst {b}")
        worst = (max(worst[0], ex), max(worst[1], eP))
    assert bk.launch_count() == 1
    if verbose:
        print(f"smoke ok: {len(ev)} events x {B} filters x {NF} features in one fused launch; "
              f"worst rel err state {worst[0]:.2e} cov {worst[1]:.2e}; results {np.bincount(res[0], minlength=5)[:5].tolist()}")
    bk.close()
    return worst

"""Generates the frozen golden vectors tests/golden/*.npz (TEST INFRASTRUCTURE).

The reference holds no golden x / P trajectory for the filter core and cannot be compiled in this image, so
these vectors are the CPU oracle's outputs (oracle/, the restatement of src/vi_ekf/*.cpp) on seeded tapes, frozen at the
commit that validated the oracle against the reference's own property tests (tests/test_oracle_properties.py).  They pin
(i) the oracle against silent drift (tests/test_golden.py::test_oracle_reproduces_golden, CPU) and (ii) the CUDA engine
against a fixture that needs neither the tape generator nor the oracle at run time (…::test_engine_matches_golden, GPU).

Run from the repo root:  python tests/golden/make_golden.py
"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

pkg = g.load_package()
abi = pkg.abi
from oracle import oracle  # noqa: E402
from tests import tapes  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

CASES = {
    # name: (NF, steps, tape seed, config overrides, instances, noise seed, bias range, perturb seed)
    "default_n12": (12, 72, 7, {}, 2, 0, None, 0),
    "default_n3": (3, 60, 5, {}, 2, 0, None, 0),
    "nodrag_nopartial_n8_noise": (8, 48, 11, dict(use_drag_term=0, use_partial_update=0, use_keyframe_reset=0), 4, 0x5EED,
                                  [0.02, 0.02, 0.02, 0.002, 0.002, 0.002], 99),
    "default_n25": (25, 40, 3, {}, 1, 0, None, 0),
}


def make(name):
    NF, steps, seed, over, count, nseed, bias, pseed = CASES[name]
    cfg = abi.default_config()
    for k, v in over.items():
        setattr(cfg, k, v)
    ev, idp, f, nb = tapes.consistent_tape(oracle, cfg, NF, steps=steps, seed=seed, extra_models=(nseed == 0))
    if nseed:
        for i in range(len(ev)):
            if ev[i]["kind"] == abi.EV_PROP:
                ev[i]["sigma"][:6] = [0.05, 0.05, 0.05, 0.002, 0.002, 0.002]
            elif ev[i]["kind"] == abi.EV_UPD and ev[i]["type"] == abi.FEAT:
                ev[i]["sigma"][:2] = 0.3
    ref = oracle.run_batch(cfg, NF, ev, idp, count=count, nthreads=2, n_buckets=nb, seed=nseed, bias_range=bias, perturb_seed=pseed)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), cfg=np.frombuffer(bytes(cfg), dtype=np.uint8), NF=NF, events=ev,
                        id_pool=np.asarray(idp, dtype=np.int32), n_buckets=nb, count=count, seed=nseed,
                        bias=np.asarray(bias if bias is not None else np.zeros(6)), perturb_seed=pseed,
                        x=ref["x"], P=ref["P"], len=ref["len"], ids=ref["ids"], results=ref["results"], stats=ref["stats"])
    print(name, "events", len(ev), "len", ref["len"].tolist(), "results", np.bincount(ref["results"][0], minlength=6)[:5].tolist())


def make_queue():
    """Streaming call pattern of BASELINE configs[0] (tests/queue_calls.py) through the oracle's rewind/replay queue."""
    from tests import queue_calls
    NF = 12
    cfg = abi.default_config()
    calls = queue_calls.config0_calls(pkg, cfg, NF)
    f = oracle.OracleVIEKF(cfg, NF)
    adds, gated, _ = queue_calls.replay(calls, f, abi, engine=False)
    os.makedirs(os.path.join(HERE, "queue"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "queue", "config0_n12.npz"), cfg=np.frombuffer(bytes(cfg), dtype=np.uint8), NF=NF, calls=calls,
                        adds=np.asarray(adds, dtype=np.int32), gated_n=np.asarray([len(g_) for g_ in gated], dtype=np.int32),
                        gated=np.asarray([i for g_ in gated for i in g_], dtype=np.int32), x=f.get_state(), P=f.get_covariance(),
                        len=f.get_len_features(), ids=np.asarray(f.tracked_features(), dtype=np.int32), t=f.time())
    print("queue/config0_n12 calls", len(calls), "adds", np.bincount(adds, minlength=5).tolist(), "gated", sum(len(g_) for g_ in gated))


if __name__ == "__main__":
    oracle.build()
    for n in (sys.argv[1:] or list(CASES) + ["queue"]):
        make_queue() if n == "queue" else make(n)

import sys, os, time, numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import __graft_entry__ as g
pkg = g.load_package()
cfg = pkg.abi.default_config()
B = int(sys.argv[1]); nimu = int(sys.argv[2]); kf = int(sys.argv[3]) if len(sys.argv) > 3 else 1
cfg.use_keyframe_reset = kf
tr = pkg.sim.simulate(cfg, 25, n_imu=nimu)
ev, idp, info = pkg.sim.build_tape(tr, cfg, 25, stat_every=20, depth_known=True)
bk = pkg.BatchVIEKF(cfg, B=B, N=25)
bk.stats_reset(info["n_buckets"])
d = bk.upload_tape(ev, idp, seed=5, bias_range=[0.5]*3+[0.05]*3)
t0 = time.time()
step = int(sys.argv[4]) if len(sys.argv) > 4 else len(ev)
for s in range(0, len(ev), step):
    bk.run_dev_tape(d, s, min(step, len(ev) - s))
    print("events", s, "..", s + step, "kinds", np.bincount(ev[s:s+step]["kind"], minlength=7).tolist(), "t", round(time.time() - t0, 2), flush=True)
print("done", time.time() - t0, bk.get_features(0, 1)[0])

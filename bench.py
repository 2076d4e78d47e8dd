#!/usr/bin/env python
"""bench.py -- filter-steps/s of the VI-EKF batch engine on B200 (BASELINE.json metric), one JSON line on rank 0.

Workload (config.workload): BASELINE.json configs[3] -- B = 262,144 filters x N = 25 features per GPU, 200 Hz IMU,
30 Hz camera with feature add/drop churn (keep_only_features every frame), drag-accelerometer update every IMU
sample, altimeter at 25 Hz (vi-ekf_b200/sim.py bench_workload; use_keyframe_reset = 0, see there).  One bench "step" = one
slice of IMU_PER_STEP IMU samples of the tape applied to every filter of the batch by ONE fused launch of the tape executor;
filter-steps = filters x IMU samples.

  value     tape already resident in HBM (viekf_batch_upload_tape), K launches timed with CUDA events on the handle's stream
  e2e       the same slices through viekf_batch_run_tape with HOST buffers: H2D of the slice's events + id pool, the
            launch, D2H of the statistics block, every step
  roofline  FP64: algorithmic flops / launch time vs the DFMA peak measured live by viekf_fp64_peak_probe (>= 60 ms kernels
            after a warm-up); `frac` counts the tape's own events (split sub-step propagates included, gated updates and no-op
            propagates excluded), `frac_8d` uses SURVEY.md 8d's fixed 656 kflop per filter-step; the HBM side is reported too
  secondary after the main timed region: BASELINE configs[1] (65,536 x 0) and configs[2] (65,536 x 8) at N = 1, the
            configs[4] shard (1,048,576 x 35 per GPU) with its statistics all-reduce at --gpus 8
  cpu_baseline / --impl reference: the CPU oracle (dense restatement of the reference, oracle/) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

METRIC = "filter-steps/sec (propagate+update, 25 features)"
UNIT = "filter-steps/s"
N_FEAT = 25
B_PER_GPU = 262144
B_PER_GPU_DEFAULT = B_PER_GPU
IMU_PER_STEP = 20
SEED = 0x5EED  # == sim.BENCH_SEED: the workload builder lives in vi-ekf_b200/sim.py so that the parity tests run this very tape
BIAS = [0.5, 0.5, 0.5, 0.05, 0.05, 0.05]  # sim_params.yaml:150-155 accel_init_stdev / gyro_init_stdev
FLOPS_8D = 656e3  # algorithmic FP64 flops per filter-step of configs[3] (366 k PROP + 3.75 FEAT + ACC + ALT / 8)


T0 = time.time()
# The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner to fd 1 at init), so the
# process's fd 1 is pointed at stderr for the whole run and the JSON line goes to a private duplicate of the original stdout.
JSON_OUT = sys.stdout


def isolate_stdout():
    global JSON_OUT
    sys.stdout.flush()
    JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def log(msg):
    if os.environ.get("VIEKF_BENCH_VERBOSE"):
        print(f"[bench {time.time() - T0:7.1f}s] {msg}", file=sys.stderr, flush=True)


def build_workload(pkg, n_slices, n_feat=None):
    return pkg.sim.bench_workload(N_FEAT if n_feat is None else n_feat, n_slices, IMU_PER_STEP)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for nme, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_leg(pkg, cfg, ev, idp, slices, n_slices, instances, threads):
    """The CPU oracle (oracle/, 'port' of the reference's dense arithmetic) over the first n_slices slices of the same
    tape, `instances` filters on `threads` host threads, one instance per thread at a time.  Returns (filter-steps/s, secs)."""
    from oracle import oracle
    end = slices[n_slices - 1][1]
    r = oracle.run_batch(cfg, N_FEAT, ev[:end], idp, count=instances, nthreads=threads, want_P=False, seed=SEED, bias_range=BIAS,
                         perturb_seed=SEED + 1, want_results=False)
    steps = instances * n_slices * IMU_PER_STEP
    return steps / r["secs"], r["secs"]


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


CPU_SAMPLE_NOTE = ("the sample starts at the feature fill-up of the first camera frame (fewer than 25 live features for the first 7 IMU "
                   "samples), so the CPU figure is slightly optimistic; the GPU arm is timed with 25 live features")


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU path.  The reference cannot be compiled in this image (Eigen3, yaml-cpp
    absent: DESIGN.md), so this is the CPU oracle port on every host thread; each step is a bounded sample of the workload."""
    if rank != 0:
        return
    pkg = g.load_package()
    from oracle import oracle
    oracle.build()
    n_slices = 2  # 40 IMU samples per instance: feature fill-up on the first frame, then 25 live features
    cfg, ev, idp, info, slices = build_workload(pkg, n_slices)
    threads = host_threads()
    inst = threads  # one instance per thread
    for _ in range(args.warmup):
        cpu_leg(pkg, cfg, ev, idp, slices, 1, inst, threads)
    t0 = time.perf_counter()
    tot_steps = 0
    for _ in range(args.steps):
        rate, secs = cpu_leg(pkg, cfg, ev, idp, slices, n_slices, inst, threads)
        tot_steps += inst * n_slices * IMU_PER_STEP
    dt = time.perf_counter() - t0
    val = tot_steps / dt
    sample = f"{inst} instances x {n_slices * IMU_PER_STEP} IMU samples of the same tape per step, {threads} threads; {CPU_SAMPLE_NOTE}"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(N_FEAT, B_PER_GPU),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=JSON_OUT, flush=True)


def kernel_name(n):
    """The executor launch_tape selects for N features (csrc/viekf_kernel.cu group_size_for, csrc/viekf_tile.cu tile_warps_for)."""
    if n == 0 and os.environ.get("VIEKF_GROUP", "8") == "8":
        return "viekf_octet_kernel (N = 0 executor: eight lanes per filter, covariance in registers)"
    if 1 <= n <= 8 and os.environ.get("VIEKF_GROUP", "2") == "2":
        return "viekf_frag_kernel (1 <= N <= 8 executor: one warp per filter, covariance in FP64 tensor-core fragments)"
    if 7 <= n <= 14:
        return "viekf_tile_kernel<1,*> (register-tile tape executor, one tile warp)"
    if 15 <= n <= 26:
        return "viekf_tile_kernel<3,3> (register-tile tape executor, 3 filters per CTA)"
    if 27 <= n <= 36:
        return "viekf_tile_kernel<6,2> (register-tile tape executor, 2 filters per CTA)"
    return "viekf_tape_kernel (shared-memory tape executor)"


def workload_config(n_feat, b_per_gpu):
    key = (n_feat, b_per_gpu)
    name = {(25, 262144): "BASELINE configs[3]", (35, 1048576): "BASELINE configs[4] per-GPU shard",
            (8, 65536): "BASELINE configs[2] (with the configs[3] event mix)",
            (0, 65536): "BASELINE configs[1] (IMU propagation + ACC-drag / ALT updates)"}.get(key, "non-default size of the configs[3] workload")
    gb = b_per_gpu * (17 + 5 * n_feat + (16 + 3 * n_feat) * (17 + 3 * n_feat) // 2) * 8 / 1e9
    return {"workload": f"{name}: {b_per_gpu} filters x {n_feat} features per GPU, 200 Hz IMU / 30 Hz camera, ACC-drag every IMU "
                        f"sample, ALT 25 Hz, keep_only_features churn, use_keyframe_reset = 0 (the altimeter measures the global altitude), "
                        f"Philox per-instance noise; step = {IMU_PER_STEP} IMU samples per filter",
            "filters_per_gpu": b_per_gpu, "features": n_feat, "imu_per_step": IMU_PER_STEP,
            "l2": f"per-step working set {gb:.2f} GB per GPU (x, P of every filter) vs 126 MB L2: "
                  + ("no flush needed" if gb > 0.5 else "fits L2; every launch reads and rewrites all of it")}


def measure(pkg, torch, dist, n_feat, B, K, W, rank, world, local_rank, want_e2e, sample_clocks):
    """W warm-up + K timed fused launches of the resident tape (value), then K steps through the host-buffer entry point (e2e).
    Returns a dict of raw measurements of this rank (times already max-reduced over ranks)."""
    n_slices = W + K + ((1 + K) if want_e2e else 0)
    cfg, ev, idp, info, slices = build_workload(pkg, n_slices, n_feat)
    nb = info["n_buckets"]
    bk = pkg.BatchVIEKF(cfg, B=B, N=n_feat, device=local_rank)
    bk.set_instance_offset(rank * B)
    bk.perturb_x0(SEED + 1)
    bk.stats_reset(nb)
    stream = torch.cuda.ExternalStream(bk.stream(), device=local_rank)
    dtape = bk.upload_tape(ev, idp, seed=SEED, bias_range=BIAS)
    log(f"N={n_feat} B={B}: workload built: {len(ev)} events, {n_slices} slices")

    def barrier():
        bk.sync()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    # ---- warm-up (includes the feature fill-up of the first camera frames) ----
    for i in range(W):
        s, e = slices[i]
        bk.run_dev_tape(dtape, s, e - s, sync=False)
    barrier()
    _, ctr0 = bk.get_flags()
    # ---- value: K fused launches, tape and state resident in HBM ----
    sampler = ClockSampler(local_rank) if sample_clocks else None
    if sampler:
        sampler.start()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    l0 = bk.launch_count()
    barrier()
    evs[0].record(stream)
    for i in range(K):
        s, e = slices[W + i]
        bk.run_dev_tape(dtape, s, e - s, sync=False)
        evs[i + 1].record(stream)
    barrier()
    clocks = sampler.stop() if sampler else None
    launches = bk.launch_count() - l0
    ms_total = evs[0].elapsed_time(evs[K])
    per_launch_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(K)]
    fl1, ctr1 = bk.get_flags()
    # ---- e2e: the same kind of slices through the host-buffer entry point, statistics read back every step ----
    ms_e2e, h2d, d2h = None, 0, 0
    if want_e2e:
        s, e = slices[W + K]
        bk.run_tape(ev[s:e], idp, seed=SEED, bias_range=BIAS)  # one untimed call: allocator / pinned staging warm
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for i in range(K):
            s, e = slices[W + K + 1 + i]
            sl = ev[s:e]
            bk.run_tape(sl, idp, seed=SEED, bias_range=BIAS)
            st = bk.stats(nb)
            h2d += sl.nbytes + idp.nbytes
            d2h += st.nbytes
        e1.record(stream)
        barrier()
        ms_e2e = e0.elapsed_time(e1)
    # ---- max over ranks, statistics all-reduce (the only collective of the path) ----
    t = torch.tensor([ms_total, ms_e2e or 0.0], dtype=torch.float64, device="cuda")
    nccl_ms = None
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)

        class _Dev:
            __cuda_array_interface__ = {"shape": (nb * pkg.abi.STAT_WIDTH,), "typestr": "<f8", "data": (bk.stats_dev_ptr(), False), "version": 3}
        stats_dev = torch.as_tensor(_Dev(), device=f"cuda:{local_rank}")
        torch.cuda.synchronize()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        dist.all_reduce(stats_dev, op=dist.ReduceOp.SUM)
        c1.record()
        torch.cuda.synchronize()
        nccl_ms = c0.elapsed_time(c1)
    ms_total, ms_e2e = float(t[0]), (float(t[1]) if want_e2e else None)
    stats = bk.stats(nb)
    bk.free_tape(dtape)
    bk.close()
    # ---- algorithmic work of the timed slices ----
    fp = fu = 0.0
    nup = 0
    for i in range(K):
        s, e = slices[W + i]
        a, b, c = pkg.sim.tape_flops(ev, info["live"], s, e - s)
        fp, fu, nup = fp + a, fu + b, nup + c
    gated_timed = (ctr1[:, 0].astype(np.int64) - ctr0[:, 0].astype(np.int64))
    gated_frac = min(1.0, float(gated_timed.mean()) / max(nup, 1))
    timed = np.concatenate([ev[slices[W + i][0]:slices[W + i][1]] for i in range(K)])
    n_upd_run = int((ev[:slices[W + K - 1][1]]["kind"] == pkg.abi.EV_UPD).sum())
    return dict(n_feat=n_feat, B=B, K=K, W=W, ms_total=ms_total, ms_e2e=ms_e2e, per_launch_ms=per_launch_ms, launches=launches, clocks=clocks,
                h2d=h2d // max(K, 1), d2h=d2h // max(K, 1), nccl_ms=nccl_ms, stats=stats, flops_per_launch=B * (fp + fu * (1.0 - gated_frac)) / K,
                gated_frac=gated_frac, timed=timed, nb=nb, cfg=cfg, ev=ev, idp=idp, slices=slices,
                nan_filters=int(((fl1 & pkg.abi.FLAG_NAN_STATE) != 0).sum()),
                mostly_gated_filters=int((ctr1[:, 0].astype(np.int64) > 0.5 * n_upd_run).sum()))


def roofline_of(pkg, m, fp64_peak, sm_mhz, peaks):
    B, K, n_feat = m["B"], m["K"], m["n_feat"]
    avg_launch_s = m["ms_total"] * 1e-3 / K  # this rank's launches back to back on one stream: event time == sum of launch durations
    achieved_tf = m["flops_per_launch"] / avg_launch_s / 1e12
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_bytes = B * pkg.sim.hbm_bytes_per_launch(n_feat)
    traffic = None
    try:  # dram__bytes_read + dram__bytes_write of one launch from the ncu --set full capture, recorded per filter
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if int(tj.get("features", 25)) == n_feat:
            traffic = tj.get("dram_bytes_per_filter_per_launch") * B
    except Exception:
        pass
    arch_peak = 148 * 64 * 2 * (sm_mhz or 1965.0) * 1e6 / 1e12  # 148 SMs x 64 DFMA lanes x 2 flop x SM clock
    r = {
        "bound": "fp64", "kernel": kernel_name(n_feat), "achieved": achieved_tf, "peak": fp64_peak, "unit": "TFLOP/s",
        "frac": achieved_tf / fp64_peak if fp64_peak > 0 else None, "traffic": traffic,
        "peak_source": "viekf_fp64_peak_probe: DFMA microbenchmark, 3 kernels of >= 60 ms after a warm-up kernel, run by this bench right after "
                       "the timed region (MEASURED_PEAKS.json has no FP64 figure)",
        "peak_architectural": arch_peak,
        "flops_per_launch": m["flops_per_launch"], "avg_launch_ms": avg_launch_s * 1e3, "gated_update_fraction": m["gated_frac"],
        "hbm": {"achieved": hbm_bytes / avg_launch_s / 1e9, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_bytes / avg_launch_s / 1e9 / hbm_peak,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650"},
    }
    if n_feat == 25:
        a8 = FLOPS_8D * B * IMU_PER_STEP / avg_launch_s / 1e12
        r["achieved_8d"] = a8
        r["frac_8d"] = a8 / fp64_peak if fp64_peak > 0 else None
        r["flops_note"] = ("frac: flops of the tape's own events (an IMU interval split by a measurement is two full propagates; gated updates and "
                           "no-op propagates excluded); frac_8d: SURVEY.md 8d's 656 kflop per filter-step")
    return r


def mc_stats_of(m):
    st = m["stats"]
    last = len(st) - 1
    n = max(st[last, 0], 1.0)
    w = min(m["W"], last)
    return {"bucket": int(last), "healthy_filters": float(st[last, 0]), "nan_filters_in_bucket": float(st[last, 38]),
            "nan_filters_this_rank": m["nan_filters"], "filters_with_over_half_updates_gated_this_rank": m["mostly_gated_filters"],
            "rmse_pos": [float(v) for v in np.sqrt(st[last, 17:20] / n)], "rmse_att": [float(v) for v in np.sqrt(st[last, 23:26] / n)],
            "nees16_mean": float(st[last, 33] / max(st[last, 37], 1.0)), "nees16_defined_for": float(st[last, 37]),
            "nees6_mean": float(st[last, 34] / max(st[last, 39], 1.0)),
            "nees16_mean_first_timed_bucket": float(st[w, 33] / max(st[w, 37], 1.0)),
            "note": "sums over the healthy filters of all ranks (all-reduced); NEES averaged over the filters where the block is positive definite"}


def run_gpu(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the engine has no CPU fallback (use --impl reference for the CPU oracle leg)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    pkg = g.load_package()
    pkg.abi.load_library()
    K, W = args.steps, max(args.warmup, 3)
    m = measure(pkg, torch, dist, N_FEAT, args.filters, K, W, rank, world, local_rank, want_e2e=True, sample_clocks=True)
    log(f"main region done: {m['ms_total']:.1f} ms for {K} launches, e2e {m['ms_e2e']:.1f} ms")
    fp64_peak = pkg.batch.fp64_peak_tflops(local_rank) if rank == 0 else 0.0
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    # ---- secondary configurations: a driver-written record of BASELINE configs[1], [2] (N = 1) and [4] (N = 8) ----
    secondary = {}
    if not args.no_secondary and args.filters == B_PER_GPU_DEFAULT and N_FEAT == 25:
        sec = []
        if world == 1:
            sec = [("configs[1]", 0, 65536, 3, 3), ("configs[2]", 8, 65536, 3, 3)]
        elif world == 8:
            sec = [("configs[4]", 35, 1048576, 2, 2)]
        for name, nf, b, k2, w2 in sec:
            try:
                m2 = measure(pkg, torch, dist, nf, b, k2, w2, rank, world, local_rank, want_e2e=False, sample_clocks=(rank == 0))
            except Exception as ex:  # a secondary line never takes the headline down
                secondary[name] = {"error": repr(ex)[:200]}
                continue
            if rank == 0:
                rf = roofline_of(pkg, m2, fp64_peak, (m2["clocks"] or {}).get("sm_mhz"), peaks)
                secondary[name] = {"workload": workload_config(nf, b)["workload"], "value": world * b * k2 * IMU_PER_STEP / (m2["ms_total"] * 1e-3),
                                   "unit": UNIT, "n_gpus": world, "steps": k2, "warmup": w2, "ms_per_step": m2["ms_total"] / k2,
                                   "kernel": rf["kernel"], "fp64_achieved_tflops": rf["achieved"], "fp64_frac": rf["frac"],
                                   "hbm_frac": rf["hbm"]["frac"], "clocks": m2["clocks"], "stats_allreduce_ms": m2["nccl_ms"],
                                   "nan_filters_this_rank": m2["nan_filters"]}
            log(f"secondary {name} done")
    if rank == 0:
        emit(args, pkg, world, m, fp64_peak, peaks, secondary)
    if world > 1:
        dist.destroy_process_group()


def emit(args, pkg, world, m, fp64_peak, peaks, secondary):
    K, W, B = m["K"], m["W"], m["B"]
    steps_total = world * B * K * IMU_PER_STEP
    value = steps_total / (m["ms_total"] * 1e-3)
    e2e = steps_total / (m["ms_e2e"] * 1e-3)
    clocks = m["clocks"]
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": m["ms_total"] / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(N_FEAT, B),
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(m["h2d"]), "d2h_bytes_per_step": int(m["d2h"]), "ms_per_step": m["ms_e2e"] / K},
        "gpu_launches": int(m["launches"]), "clocks": clocks, "roofline": roofline_of(pkg, m, fp64_peak, (clocks or {}).get("sm_mhz"), peaks),
        "per_launch_ms": [round(v, 3) for v in m["per_launch_ms"]],
    }
    # per-kind event rates of the whole job over the timed region (propagates/s, FEAT updates/s, ...)
    abi = pkg.abi
    timed = m["timed"]
    upd = timed[timed["kind"] == abi.EV_UPD]
    per_s = world * B / (m["ms_total"] * 1e-3)
    line["event_rates_per_s"] = {
        "propagate": float((timed["kind"] == abi.EV_PROP).sum() * per_s),
        "update_feat": float((upd["type"] == abi.FEAT).sum() * per_s), "update_acc": float((upd["type"] == abi.ACC).sum() * per_s),
        "update_alt": float((upd["type"] == abi.ALT).sum() * per_s),
        "keep_only_features": float((timed["kind"] == abi.EV_KEEP).sum() * per_s), "init_feature": float((timed["kind"] == abi.EV_INIT).sum() * per_s),
    }
    if m["nccl_ms"] is not None:
        line["stats_allreduce_ms"] = m["nccl_ms"]
    line["mc_stats"] = mc_stats_of(m)
    if secondary:
        line["secondary"] = secondary
    if world == 1 and not args.no_cpu:
        from oracle import oracle
        oracle.build()
        threads = host_threads()
        cfg, ev, idp, slices = m["cfg"], m["ev"], m["idp"], m["slices"]
        rate0, _ = cpu_leg(pkg, cfg, ev, idp, slices, 1, threads, threads)  # calibration: one slice per thread
        inst = int(min(max(threads, round(rate0 * 15.0 / (2 * IMU_PER_STEP) / threads) * threads), 64 * threads))
        rate, secs = cpu_leg(pkg, cfg, ev, idp, slices, 2, inst, threads)
        line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"{inst} instances x {2 * IMU_PER_STEP} IMU samples (first two slices of the same tape), {secs:.1f} s; {CPU_SAMPLE_NOTE}"}
    print(json.dumps(line), file=JSON_OUT, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--filters", type=int, default=B_PER_GPU, help="filters per GPU (default: BASELINE configs[3])")
    ap.add_argument("--features", type=int, default=N_FEAT, help="features per filter (default 25: BASELINE configs[3]; 35 = configs[4])")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-secondary", action="store_true", help="skip the secondary configurations (configs[1], [2]; configs[4] at --gpus 8)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "b200" and args.gpus > 1 and world == 1 and "RANK" not in os.environ:
        # `python bench.py --gpus N` outside torchrun: launch the N ranks ourselves, exactly as the driver does
        port = 29500 + os.getpid() % 2000
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1",
               "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    isolate_stdout()
    globals().update(N_FEAT=args.features, B_PER_GPU=args.filters, METRIC=METRIC.replace("25 features", f"{args.features} features"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    g.build() if not os.path.exists(g.load_package().abi.lib_path()) else None
    run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()

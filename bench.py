#!/usr/bin/env python
"""bench.py -- filter-steps/s of the VI-EKF batch engine on B200 (BASELINE.json metric), one JSON line on rank 0.

Workload (config.workload): BASELINE.json configs[3] -- B = 262,144 filters x N = 25 features per GPU, 200 Hz IMU,
30 Hz camera with feature add/drop churn (keep_only_features every frame), drag-accelerometer update every IMU
sample, altimeter at 25 Hz.  One bench "step" = one slice of IMU_PER_STEP IMU samples of the tape applied to every
filter of the batch by ONE fused launch of the tape executor; filter-steps = filters x IMU samples.

  value     tape already resident in HBM (viekf_batch_upload_tape), K launches timed with CUDA events on the handle's stream
  e2e       the same slices through viekf_batch_run_tape with HOST buffers: H2D of the slice's events + id pool, the
            launch, D2H of the statistics block, every step
  roofline  FP64: algorithmic flops (SURVEY.md 8d model, gated updates excluded) / launch time vs the DFMA peak
            measured live by viekf_fp64_peak_probe; the HBM side of the same launch is reported next to it
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
IMU_PER_STEP = 20
SEED = 0x5EED
BIAS = [0.5, 0.5, 0.5, 0.05, 0.05, 0.05]  # sim_params.yaml:150-155 accel_init_stdev / gyro_init_stdev


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


def build_workload(pkg, n_steps_total):
    cfg = pkg.abi.default_config()
    truth = pkg.sim.simulate(cfg, N_FEAT, n_imu=n_steps_total * IMU_PER_STEP)
    ev, idp, info = pkg.sim.build_tape(truth, cfg, N_FEAT, stat_every=IMU_PER_STEP, depth_known=True)
    # slice boundaries: a slice ends with the STAT event that closes its IMU_PER_STEP-th IMU sample
    ends = [i + 1 for i in range(len(ev)) if ev[i]["kind"] == pkg.abi.EV_STAT]
    starts = [0] + ends[:-1]
    return cfg, ev, idp, info, list(zip(starts, ends))


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
    sample = f"{inst} instances x {n_slices * IMU_PER_STEP} IMU samples of the same tape per step, {threads} threads"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * dt / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=JSON_OUT, flush=True)


def kernel_name(n):
    """The executor launch_tape selects for N features (csrc/viekf_kernel.cu group_size_for, csrc/viekf_tile.cu tile_warps_for)."""
    if 7 <= n <= 14:
        return "viekf_tile_kernel<1,6> (register-tile tape executor, one tile warp, 6 filters per CTA)"
    if 15 <= n <= 26:
        return "viekf_tile_kernel<3,3> (register-tile tape executor, 3 filters per CTA)"
    if 27 <= n <= 36:
        return "viekf_tile_kernel<6,2> (register-tile tape executor, 2 filters per CTA)"
    return "viekf_tape_kernel (shared-memory tape executor)"


def workload_config():
    name = "BASELINE configs[3]" if (N_FEAT, B_PER_GPU) == (25, 262144) else (
        "BASELINE configs[4] per-GPU shard" if (N_FEAT, B_PER_GPU) == (35, 1048576) else (
        "BASELINE configs[2] (with the configs[3] event mix)" if (N_FEAT, B_PER_GPU) == (8, 65536) else (
        "BASELINE configs[1] (IMU propagation + ACC-drag / ALT updates)" if (N_FEAT, B_PER_GPU) == (0, 65536) else "non-default size of the configs[3] workload")))
    gb = B_PER_GPU * (17 + 5 * N_FEAT + (16 + 3 * N_FEAT) * (17 + 3 * N_FEAT) // 2) * 8 / 1e9
    return {"workload": f"{name}: {B_PER_GPU} filters x {N_FEAT} features per GPU, 200 Hz IMU / 30 Hz camera, ACC-drag every IMU "
                        f"sample, ALT 25 Hz, keep_only_features churn, Philox per-instance noise; step = {IMU_PER_STEP} IMU samples per filter",
            "filters_per_gpu": B_PER_GPU, "features": N_FEAT, "imu_per_step": IMU_PER_STEP,
            "l2": f"per-step working set {gb:.1f} GB per GPU (x, P of every filter) >> 126 MB L2, no flush needed"}


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
    n_slices = W + K + 1 + K
    cfg, ev, idp, info, slices = build_workload(pkg, n_slices)
    nb = info["n_buckets"]
    B = args.filters
    bk = pkg.BatchVIEKF(cfg, B=B, N=N_FEAT, device=local_rank)
    bk.set_instance_offset(rank * B)
    bk.perturb_x0(SEED + 1)
    bk.stats_reset(nb)
    fp64_peak = pkg.batch.fp64_peak_tflops(local_rank)
    stream = torch.cuda.ExternalStream(bk.stream(), device=local_rank)
    dtape = bk.upload_tape(ev, idp, seed=SEED, bias_range=BIAS)
    log(f"workload built: {len(ev)} events, {n_slices} slices, fp64 peak {fp64_peak:.1f} TF")

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
    log("warm-up done")
    _, ctr0 = bk.get_flags(0, min(B, 4096))
    # ---- value: K fused launches, tape and state resident in HBM ----
    sampler = ClockSampler(local_rank)
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
    clocks = sampler.stop()
    launches = bk.launch_count() - l0
    ms_total = evs[0].elapsed_time(evs[K])
    log(f"value region done: {ms_total:.1f} ms for {K} launches")
    per_launch_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(K)]
    _, ctr1 = bk.get_flags(0, min(B, 4096))
    # ---- e2e: the same kind of slices through the host-buffer entry point, statistics read back every step ----
    s, e = slices[W + K]
    bk.run_tape(ev[s:e], idp, seed=SEED, bias_range=BIAS)  # one untimed call: allocator / pinned staging warm
    h2d = d2h = 0
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
    log(f"e2e region done: {ms_e2e:.1f} ms")
    # ---- max over ranks, statistics all-reduce (the only collective of the path) ----
    t = torch.tensor([ms_total, ms_e2e], dtype=torch.float64, device="cuda")
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
    ms_total, ms_e2e = float(t[0]), float(t[1])
    stats = bk.stats(nb)
    if rank == 0:
        # the roofline denominator: the better of the probe before the run (cold clocks) and after it (boost clocks held)
        fp64_peak = max(fp64_peak, pkg.batch.fp64_peak_tflops(local_rank))
        emit(args, pkg, world, K, W, B, ev, idp, info, slices, cfg, ms_total, ms_e2e, per_launch_ms, launches, clocks, fp64_peak, ctr0, ctr1, stats,
             h2d // K, d2h // K, nccl_ms)
    bk.free_tape(dtape)
    bk.close()
    if world > 1:
        dist.destroy_process_group()


def emit(args, pkg, world, K, W, B, ev, idp, info, slices, cfg, ms_total, ms_e2e, per_launch_ms, launches, clocks, fp64_peak, ctr0, ctr1, stats,
         h2d, d2h, nccl_ms):
    steps_total = world * B * K * IMU_PER_STEP
    value = steps_total / (ms_total * 1e-3)
    e2e = steps_total / (ms_e2e * 1e-3)
    # ---- roofline of the dominant (only) kernel: viekf_tape_kernel ----
    fp = fu = 0.0
    nup = 0
    for i in range(K):
        s, e = slices[W + i]
        a, b, c = pkg.sim.tape_flops(ev, info["live"], s, e - s)
        fp, fu, nup = fp + a, fu + b, nup + c
    gated = float((ctr1[:, 0].astype(np.int64) - ctr0[:, 0].astype(np.int64)).mean())  # gated updates per filter in the timed region
    gated_frac = min(1.0, gated / max(nup, 1))
    flops_per_launch = B * (fp + fu * (1.0 - gated_frac)) / K
    avg_launch_s = ms_total * 1e-3 / K  # this rank's launches back to back on one stream: event time == sum of launch durations
    achieved_tf = flops_per_launch / avg_launch_s / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_bytes = B * pkg.sim.hbm_bytes_per_launch(N_FEAT)
    traffic = None
    try:  # dram__bytes_read + dram__bytes_write of one launch from the ncu --set full capture, recorded per filter
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("dram_bytes_per_filter_per_launch") * B
    except Exception:
        pass
    roofline = {
        "bound": "fp64", "kernel": kernel_name(N_FEAT), "achieved": achieved_tf, "peak": fp64_peak, "unit": "TFLOP/s",
        "frac": achieved_tf / fp64_peak if fp64_peak > 0 else None, "traffic": traffic,
        "peak_source": "viekf_fp64_peak_probe (DFMA microbenchmark run by this bench; MEASURED_PEAKS.json has no FP64 figure)",
        "flops_per_launch": flops_per_launch, "avg_launch_ms": avg_launch_s * 1e3, "gated_update_fraction": gated_frac,
        "hbm": {"achieved": hbm_bytes / avg_launch_s / 1e9, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_bytes / avg_launch_s / 1e9 / hbm_peak,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650"},
    }
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_total / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(),
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e / K},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
        "per_launch_ms": [round(v, 3) for v in per_launch_ms],
    }
    line["config"]["filters_per_gpu"] = B
    # per-kind event rates of the whole job over the timed region (propagates/s, FEAT updates/s, ...)
    abi = pkg.abi
    timed = np.concatenate([ev[slices[W + i][0]:slices[W + i][1]] for i in range(K)])
    upd = timed[timed["kind"] == abi.EV_UPD]
    per_s = world * B / (ms_total * 1e-3)
    line["event_rates_per_s"] = {
        "propagate": float((timed["kind"] == abi.EV_PROP).sum() * per_s),
        "update_feat": float((upd["type"] == abi.FEAT).sum() * per_s), "update_acc": float((upd["type"] == abi.ACC).sum() * per_s),
        "update_alt": float((upd["type"] == abi.ALT).sum() * per_s),
        "keep_only_features": float((timed["kind"] == abi.EV_KEEP).sum() * per_s), "init_feature": float((timed["kind"] == abi.EV_INIT).sum() * per_s),
    }
    if nccl_ms is not None:
        line["stats_allreduce_ms"] = nccl_ms
    n = np.maximum(stats[:, 0], 1)
    last = -1
    line["mc_stats"] = {"count": float(stats[last, 0]), "rmse_pos": [float(v) for v in np.sqrt(stats[last, 17:20] / n[last])],
                        "rmse_att": [float(v) for v in np.sqrt(stats[last, 23:26] / n[last])], "nees16_mean": float(stats[last, 33] / n[last])}
    if world == 1 and not args.no_cpu:
        from oracle import oracle
        oracle.build()
        threads = host_threads()
        rate0, _ = cpu_leg(pkg, cfg, ev, idp, slices, 1, threads, threads)  # calibration: one slice per thread
        inst = int(min(max(threads, round(rate0 * 15.0 / (2 * IMU_PER_STEP) / threads) * threads), 64 * threads))
        rate, secs = cpu_leg(pkg, cfg, ev, idp, slices, 2, inst, threads)
        line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"{inst} instances x {2 * IMU_PER_STEP} IMU samples (first two slices of the same tape), {secs:.1f} s"}
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
    args = ap.parse_args()
    isolate_stdout()
    globals().update(N_FEAT=args.features, B_PER_GPU=args.filters, METRIC=METRIC.replace("25 features", f"{args.features} features"))
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    g.build() if not os.path.exists(g.load_package().abi.lib_path()) else None
    run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()

"""Phase timing of the tile executor (tuning helper).  Needs a build with -DVIEKF_PHASE_TIMING (make -C vi-ekf_b200/csrc pht;
VIEKF_LIB=.../libviekf_b200_pht.so): runs bench.py's workload, then prints the elapsed clocks per phase and warp role
(tile warps, base warp, x warp; see PH() in viekf_tile.cu)."""
import ctypes, os, sys, io, json
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import bench  # noqa: E402
sys.argv = ["bench.py", "--no-cpu", "--no-secondary", "--filters", sys.argv[1] if len(sys.argv) > 1 else "32768", "--steps", "3", "--warmup", "3"]
buf = io.StringIO()
bench.isolate_stdout = lambda: None  # keep the JSON line in this process
bench.JSON_OUT = buf
bench.main()
d = json.loads([l for l in buf.getvalue().splitlines() if l.startswith("{")][-1])
print("value %.4g ms/launch %s" % (d["value"], d["per_launch_ms"]))
lib = ctypes.CDLL(bench.g.load_package().abi.lib_path())
NR = 8
clk = (ctypes.c_ulonglong * (NR * 64))(); cnt = (ctypes.c_ulonglong * (NR * 64))()
lib.viekf_debug_phase_clocks(clk, cnt, 0)
names = {0: "load", 1: "PROP wait model/phase A", 2: "PROP B rows", 3: "PROP D1 | C,P'bb | Y,boxplus", 4: "PROP D2 | P'fb | resolve", 8: "UPD head", 9: "UPD P1 PHt | resolve", 10: "UPD P2 gate (x)",
         11: "UPD P3 K,W rows", 12: "UPD P4 x: boxplus+fix", 13: "UPD P4 tile rank", 14: "UPD P4 base rank",
         15: "store", 16: "end PROP (x: prepare)", 17: "end UPD (x: prepare)", 18: "INIT", 19: "CLEAR", 20: "KEEP", 21: "STAT", 22: "KFRESET"}
roles = [w for w in range(NR) if any(cnt[64 * w + k] for k in range(64))]
tot = sum(clk[64 * roles[0] + k] for k in range(64))
print("phase: per warp role (" + ", ".join(f"w{w}" for w in roles) + ")  work clk / wait-at-barrier clk per occurrence   (share of warp 0's time)")
for k in range(32):
    if any(cnt[64 * w + k] for w in roles):
        cells = []
        for w in roles:
            wk = clk[64 * w + k] / max(cnt[64 * w + k], 1)
            wt = clk[64 * w + k + 32] / max(cnt[64 * w + k + 32], 1)
            cells.append(f"{wk:7.0f}/{wt:6.0f}")
        w0 = roles[0]
        print(f"{names.get(k, k):30s} " + "  ".join(cells) + f"   {100.0 * (clk[64 * w0 + k] + clk[64 * w0 + k + 32]) / tot:5.1f}%  x{cnt[64 * w0 + k]}")

"""Phase timing of the tile executor (tuning helper).  Needs a build with NVFLAGS+=-DVIEKF_PHASE_TIMING:
runs bench.py's workload, then prints the elapsed clocks per phase on the CTAs' critical path (see PH() in viekf_tile.cu)."""
import ctypes, os, sys, io, contextlib, json
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
clk = (ctypes.c_ulonglong * 256)(); cnt = (ctypes.c_ulonglong * 256)()
lib.viekf_debug_phase_clocks(clk, cnt, 0)
names = {0: "load", 1: "PROP head+A+B", 2: "PROP C", 3: "PROP D1", 4: "PROP D2", 8: "UPD head+H", 9: "UPD PHt", 10: "UPD S,inv,K", 11: "UPD resolve+rest", 12: "UPD  boxplus", 13: "UPD  tile rank", 14: "UPD  base rank",
         15: "store", 16: "PROP F", 17: "UPD fix (or gated/new)", 18: "INIT", 19: "CLEAR", 20: "KEEP", 21: "STAT", 22: "KFRESET"}
tot = sum(clk[0:64])
print("phase: per warp  work clk / wait-at-barrier clk per occurrence   (share of warp 0's time)")
for k in range(32):
    if cnt[k]:
        cells = []
        for w in range(4):
            wk = clk[64 * w + k] / max(cnt[64 * w + k], 1)
            wt = clk[64 * w + k + 32] / max(cnt[64 * w + k + 32], 1)
            cells.append(f"{wk:7.0f}/{wt:6.0f}")
        print(f"{names.get(k, k):26s} " + "  ".join(cells) + f"   {100.0 * (clk[k] + clk[k + 32]) / tot:5.1f}%  x{cnt[k]}")

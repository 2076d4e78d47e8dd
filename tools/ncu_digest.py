"""Digest of an .ncu-rep: key raw metrics + top CUDA source lines by stall samples (tuning helper)."""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
keep = ['gpu__time_duration.sum','launch__registers_per_thread','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','launch__grid_size','launch__block_size','sm__warps_active.avg.pct_of_peak_sustained_active','dram__bytes_read.sum','dram__bytes_write.sum','smsp__inst_executed.sum','sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio','smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio','smsp__average_warps_issue_stalled_wait_per_issue_active.ratio','smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio','smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio','smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio','smsp__average_warp_latency_per_inst_issued.ratio','smsp__sass_inst_executed_op_local_ld.sum']
print("metric,unit,value")
for h, u, v in zip(hdr, units, vals):
    if h in keep: print(f"{h},{u},{v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
cur = None; hd = None; out = []
for r in rows:
    if len(r) == 2 and r[0] == 'File Path': cur = r[1]; continue
    if r and r[0] == 'Line No': hd = r; continue
    if hd and len(r) == len(hd) and r[0].isdigit():
        d = dict(zip(hd, r))
        g = lambda k: int(d.get(k) or 0)
        out.append((cur.split('/')[-1], int(r[0]), r[1], g('# Samples'), g('Instructions Executed'), g('stall_barrier'), g('stall_short_sb'), g('L1 Wavefronts Shared'), g('L1 Wavefronts Shared Excessive')))
ts = sum(o[3] for o in out) or 1; ti = sum(o[4] for o in out) or 1; tw = sum(o[7] for o in out) or 1
print(f"# total samples {ts} inst {ti} shared wavefronts {tw}")
print("# file:line samples% inst% barrier short_sb smem_wf% excessive_wf | source")
for o in sorted(out, key=lambda o: -o[3])[:top]:
    print(f"{o[0]}:{o[1]} {100*o[3]/ts:.1f}% {100*o[4]/ti:.1f}% {o[5]} {o[6]} {100*o[7]/tw:.1f}% {o[8]} | {o[2].strip()[:100]}")

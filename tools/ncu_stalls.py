"""Prints every warp-stall reason (cycles per issued instruction) and a few throughput metrics of an .ncu-rep (tuning helper)."""
import csv, subprocess, sys, io
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); h, u, v = rows[0], rows[1], rows[2]
for a, b, c in zip(h, u, v):
    if 'issue_stalled' in a and 'per_issue_active' in a and 'not_issued' not in a and float(c) > 0.05:
        print(a.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''), c)
    if a in ('gpu__time_duration.sum', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
             'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__sass_inst_executed_op_local_ld.sum', 'smsp__sass_inst_executed_op_local_st.sum',
             'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum'):
        print(a, c, b)

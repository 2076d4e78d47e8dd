"""Aggregates an .ncu-rep source page by file and line bucket (tuning helper): samples%, inst%, barrier-stall samples."""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]; bucket = int(sys.argv[2]) if len(sys.argv) > 2 else 25
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
cur = None; hd = None
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
for r in rows:
    if len(r) == 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r and r[0] == 'Line No': hd = r; continue
    if hd and len(r) == len(hd) and r[0].isdigit():
        d = dict(zip(hd, r)); g = lambda k: int(d.get(k) or 0)
        key = (cur, (int(r[0]) // bucket) * bucket)
        a = agg[key]; a[0] += g('# Samples'); a[1] += g('Instructions Executed'); a[2] += g('stall_barrier'); a[3] += g('L1 Wavefronts Shared')
ts = sum(a[0] for a in agg.values()) or 1; ti = sum(a[1] for a in agg.values()) or 1; tw = sum(a[3] for a in agg.values()) or 1
print("file:line-bucket samples% inst% barrier% smem_wf%")
for k, a in sorted(agg.items()):
    if a[0] * 200 > ts or a[1] * 200 > ti:
        print(f"{k[0]}:{k[1]} {100*a[0]/ts:.1f}% {100*a[1]/ti:.1f}% {100*a[2]/ts:.1f}% {100*a[3]/tw:.1f}%")

"""Lists the SASS local-memory loads/stores of an .ncu-rep by executed count (tuning helper)."""
import csv, subprocess, sys, io
rep = sys.argv[1]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hd = None; data = []
for r in rows:
    if r and r[0] == 'Address': hd = r; continue
    if hd and len(r) == len(hd): data.append(dict(zip(hd, r)))
tot = sum(int(d['Instructions Executed'] or 0) for d in data)
loc = [(i, d) for i, d in enumerate(data) if 'LDL' in d['Source'] or 'STL' in d['Source']]
print("total inst", tot, "local inst", sum(int(d['Instructions Executed'] or 0) for _, d in loc), "static local", len(loc), "of", len(data))
for i, d in sorted(loc, key=lambda t: -int(t[1]['Instructions Executed'] or 0))[:40]:
    print(i, d['Instructions Executed'], d['# Samples'], d['Source'][:70])

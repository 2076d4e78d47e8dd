"""Static SASS instruction count per source file / line bucket from `nvdisasm -g -c` output (tuning helper)."""
import re, collections, sys
path = sys.argv[1]; bucket = int(sys.argv[2]) if len(sys.argv) > 2 else 25
cnt = collections.Counter(); cur = None; fn = None; per_fn = collections.Counter()
for line in open(path):
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2)) // bucket * bucket); continue
    m = re.match(r'\s*\.text\.(\S+):', line)
    if m: fn = m.group(1)[:60]
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', line):
        cnt[(fn, cur)] += 1; per_fn[fn] += 1
for f, v in per_fn.items(): print("FUNCTION", f, v)
want = sys.argv[3] if len(sys.argv) > 3 else None
for (f, k), v in sorted(cnt.items(), key=lambda kv: -kv[1])[:70]:
    if want is None or want in (f or ''): print(f[:40] if f else f, k, v)

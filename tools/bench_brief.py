"""Runs bench.py with the given extra args and prints a one-line digest (tuning helper, not part of the product)."""
import json, subprocess, sys, os
args = sys.argv[1:]
p = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "..", "bench.py")] + args, capture_output=True, text=True, timeout=200)
lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
if not lines:
    print("NO JSON; rc", p.returncode, p.stderr[-800:])
    sys.exit(0)
d = json.loads(lines[-1])
print(f"G={os.environ.get('VIEKF_GROUP','auto')} value={d['value']:.4g} e2e={d['e2e']['value']:.4g} ms/launch={d['per_launch_ms']} frac={d['roofline']['frac']:.4f} "
      f"clk={d['clocks'].get('sm_mhz')} cpu={d.get('cpu_baseline',{}).get('value')}")

"""Manual helper: run the default bench N times in fresh processes and print one short line each."""
import json, subprocess, sys
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
extra = sys.argv[2:]
for i in range(n):
    out = subprocess.run([sys.executable, "bench.py", *extra], capture_output=True, text=True).stdout.strip().splitlines()
    l = json.loads(out[-1]); r = l["roofline"]
    e = l.get("e2e", {})
    print(f"run {i}: {l['value']:.0f} fits/s {l['ms_per_step']:.2f} ms warm {l['warmup']} frac {r['frac']:.3f} kernel {r['kernel_ms_per_step']:.2f} ms "
          f"share {r['kernel_share_of_step']:.3f} sm {l['clocks']['sm_mhz']} e2e {e.get('value')}", flush=True)

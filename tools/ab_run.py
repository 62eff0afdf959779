"""Manual A/B helper (not a pytest file): python tools/ab_run.py STEPS VAR=a,b [VAR2=c,d] ...
Runs bench.py --no-e2e --no-cpu-baseline once per value (interleaved twice) and prints one line each."""
import json, os, subprocess, sys
steps = sys.argv[1]
var, vals = sys.argv[2].split("=")
for rep in range(2):
    for v in vals.split(","):
        env = dict(os.environ, **{var: v})
        out = subprocess.run([sys.executable, "bench.py", "--no-e2e", "--no-cpu-baseline", "--steps", steps],
                             env=env, capture_output=True, text=True).stdout.strip().splitlines()
        l = json.loads(out[-1])
        print(f"{var}={v}: {l['value']:.0f} fits/s {l['ms_per_step']:.3f} ms frac {l['roofline']['frac']:.3f} "
              f"sm {l['clocks']['sm_mhz']} MHz {l['clocks']['power_w']} W", flush=True)

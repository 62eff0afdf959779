"""Manual A/B helper (not a pytest file): python tools/ab_lib.py STEPS LIB_A.so LIB_B.so ...
Compares builds of the library: copies each file over libdml_b200.so in turn (interleaved twice), runs
bench.py --no-e2e --no-cpu-baseline and prints one line per run; the last file stays in place."""
import json, os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cs230_distributed_machine_learning_b200", "libdml_b200.so")
steps, libs = sys.argv[1], sys.argv[2:]
for rep in range(2):
    for lib in libs:
        shutil.copyfile(lib, LIB + ".tmp"); os.replace(LIB + ".tmp", LIB)
        out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--no-e2e", "--no-cpu-baseline", "--steps", steps],
                             capture_output=True, text=True).stdout.strip().splitlines()
        l = json.loads(out[-1])
        print(f"{os.path.basename(lib)}: {l['value']:.0f} fits/s {l['ms_per_step']:.3f} ms frac {l['roofline']['frac']:.3f} "
              f"sm {l['clocks']['sm_mhz']} MHz {l['clocks']['power_w']} W", flush=True)

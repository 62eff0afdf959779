"""Opcode histogram per kernel of the in-tree library (cuobjdump -sass): which kernels issue tcgen05 MMAs
(UTCHMMA / UTCQMMA ...), tensor-memory loads and stores (LDTM / STTM), TMA loads (UTMALDG) and multicast.

    python tools/sass_summary.py > profiles/r02_sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cs230_distributed_machine_learning_b200", "libdml_b200.so")
WATCH = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTCCP", "LDTM", "STTM", "UTMALDG", "UTMAPF", "SYNCS", "UCGABAR", "FFMA", "DFMA", "MUFU"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
    kernels, cur = collections.OrderedDict(), None
    it = iter(names)
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = next(it, m.group(1))
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur is not None:
            op = m.group(1)
            kernels[cur]["_total"] += 1
            for w in WATCH:
                if op.startswith(w):
                    kernels[cur][w] += 1
            if op.startswith("UTMALDG") and "MULTICAST" in op:
                kernels[cur]["UTMALDG.MULTICAST"] += 1
            if op.startswith("UTCHMMA") and "2CTA" in op:
                kernels[cur]["UTCHMMA.2CTA"] += 1
    cols = WATCH + ["UTMALDG.MULTICAST", "UTCHMMA.2CTA"]
    print(f"# {os.path.relpath(LIB, ROOT)}: SASS opcode counts per kernel (cuobjdump -sass, sm_100a)")
    print("# columns: total instructions, then " + ", ".join(cols))
    tot = collections.Counter()
    for k, c in kernels.items():
        depth, cut = 0, len(k)
        for i, ch in enumerate(k):          # cut the parameter list: the first "(" outside the template arguments
            depth += (ch == "<") - (ch == ">")
            if ch == "(" and depth == 0:
                cut = i
                break
        short = k[:cut].replace("(int)", "").replace("void ", "")
        print(f"{short[:110]:110s} {c['_total']:7d} " + " ".join(f"{c[w]:6d}" for w in cols))
        tot.update(c)
    print(f"{'ALL KERNELS':110s} {tot['_total']:7d} " + " ".join(f"{tot[w]:6d}" for w in cols))


if __name__ == "__main__":
    sys.exit(main())

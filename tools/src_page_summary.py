"""Condense an `ncu --page source --csv` export into a text summary (stall mix, hot segments, top instructions).

    python tools/src_page_summary.py gpurun_out/a_src_fwd_cfg4.csv [section] > profiles/r02_ncu_src_fwd_cfg4_after.txt

A segment = a run of consecutive SASS instructions with the same execution count (i.e. one loop level).
"""
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    secs = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
    a = secs[int(sys.argv[2]) if len(sys.argv) > 2 else 0]
    later = [s for s in secs if s > a]
    b = later[0] if later else len(rows)
    print("kernel:", rows[a][1])
    hdr = rows[a + 1]
    data = [r for r in rows[a + 2:b] if len(r) == len(hdr)]
    ix = {h: i for i, h in enumerate(hdr)}
    S = lambda r: int(r[ix["# Samples"]])
    E = lambda r: int(r[ix["Instructions Executed"]])
    tot = sum(S(r) for r in data)
    print(f"SASS instructions: {len(data)}   warp samples: {tot}   warp instructions executed: {sum(E(r) for r in data)}")
    st = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = {s: sum(int(r[ix[s]]) for r in data) for s in st}
    print("\nstall mix (all samples):")
    for k, v in sorted(agg.items(), key=lambda x: -x[1])[:10]:
        print(f"  {k[6:]:<20} {v:>9}  {v / tot:6.1%}")
    seg = []
    for i, r in enumerate(data):
        e = E(r)
        if seg and seg[-1][2] == e:
            seg[-1][1] = i; seg[-1][3] += S(r)
        else:
            seg.append([i, i, e, S(r)])
    print("\nsegments with >= 1 % of the samples (first..last instruction, executions per instruction, samples, share):")
    for x, y, e, s in seg:
        if s >= tot * 0.01:
            print(f"  {x:>5}..{y:<5} n={y - x + 1:<4} exec={e:<11} samples={s:<9} {s / tot:6.1%}")
    by_exec = {}
    for r in data:
        by_exec.setdefault(E(r), [0, 0])
        by_exec[E(r)][0] += 1; by_exec[E(r)][1] += S(r)
    print("\ninstructions per execution count (count, instructions, samples):")
    for e, (n, s) in sorted(by_exec.items(), key=lambda x: -x[1][1])[:8]:
        print(f"  exec={e:<11} instructions={n:<5} samples={s:<9} {s / tot:6.1%}")
    print("\ntop 30 instructions by samples (index, samples, executions, SASS, main stall, L1 tag requests):")
    for i in sorted(range(len(data)), key=lambda i: -S(data[i]))[:30]:
        r = data[i]
        tp = max(((int(r[ix[h]]), h[6:]) for h in st))
        tags = r[ix["L1 Tag Requests Global"]] if "L1 Tag Requests Global" in ix else "-"
        print(f"  {i:>5} {S(r):>8} {E(r):>11}  {r[1].strip()[:72]:<72} {tp[1]}:{tp[0]}  tags={tags}")


if __name__ == "__main__":
    main()

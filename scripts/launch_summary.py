"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel totals and shares."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
h, data = rows[hdr], rows[hdr + 1:]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
gi = h.index("Grid Size") if "Grid Size" in h else None
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
seq = []
for r in data:
    v = float(r[vi].replace(",", ""))
    u = r[ui]
    v = v / 1e6 if u.startswith("n") else (v / 1e3 if u.startswith("u") else v)
    name = r[ki].split("(")[0].replace("void ", "")[:70]
    agg[name][0] += 1
    agg[name][1] += v
    tot += v
    seq.append((name, v, r[gi] if gi is not None else ""))
print(f"total {tot:.3f} ms over {len(data)} launches (cold-cache, serialised: compare shares)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[1]:9.3f} ms {100 * v[1] / tot:5.1f}%  n={v[0]:4d}  {k}")
if len(sys.argv) > 2:
    print("--- per launch")
    for i, (n, v, g) in enumerate(seq):
        print(f"{i:4d} {v:8.4f} ms grid={g:>14s} {n}")

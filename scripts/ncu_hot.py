"""Top stall-sample SASS lines of one kernel in an .ncu-rep source page export (csv)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]
data = [r for r in rows[2:] if len(r) > 5]
si, ie = h.index("Warp Stall Sampling (All Samples)"), h.index("Instructions Executed")
num = lambda x: int(x) if x.strip().isdigit() else 0
tot = sum(num(r[si]) for r in data)
print("total samples", tot, "warp instructions", sum(num(r[ie]) for r in data))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
for r in sorted(data, key=lambda r: -num(r[si]))[:n]:
    print(r[si].rjust(6), f"{100 * num(r[si]) / tot:5.1f}%", r[ie].rjust(9), r[1][:100])

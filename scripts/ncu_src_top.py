"""Top stall lines of an ncu report's source page: python scripts/ncu_src_top.py report.ncu-rep [N]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = rows[1]; idx = {k: i for i, k in enumerate(h)}; data = rows[2:]
tot = sum(int(r[idx['# Samples']]) for r in data)
stalls = [k for k in h if k.startswith('stall_') and 'Not Issued' not in k]
agg = {k: sum(int(r[idx[k]]) for r in data) for k in stalls}
print(rows[0][1][:80], 'total samples', tot)
print('by reason:', sorted(((v, k) for k, v in agg.items() if v), reverse=True)[:8])
for r in sorted(data, key=lambda r: -int(r[idx['# Samples']]))[:N]:
    s = int(r[idx['# Samples']])
    st = sorted(((int(r[idx[k]]), k[6:]) for k in stalls), reverse=True)[:2]
    print(f"{s:6d} {100*s/tot:5.1f}% x{r[idx['Instructions Executed']]:>8s} {r[idx['Source']].strip()[:64]:64s} {st}")

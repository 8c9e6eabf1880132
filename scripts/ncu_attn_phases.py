"""Stall samples of attn_fwd_v2_kernel bucketed by phase of the softmax loop, from the source page of an
`ncu --set full --import-source on` report:  python scripts/ncu_attn_phases.py report.ncu-rep"""
import collections
import csv
import io
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = rows[1]
idx = {k: i for i, k in enumerate(h)}
data = rows[2:]
stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
src = [r[idx["Source"]].strip() for r in data]
S = lambda r: int(r[idx["# Samples"]])


def first(pred, start=0):
    return next(i for i in range(start, len(data)) if pred(src[i]))


ldtm = first(lambda s: s.startswith("LDTM"))
wait_s = max(i for i in range(ldtm) if "TRYWAIT" in src[i])
arrive = first(lambda s: "SYNCS.ARRIVE" in s, ldtm)
fmn = first(lambda s: s.startswith("FMNMX3"), arrive)
ffma = first(lambda s: s.startswith("FFMA2"), fmn)
last_f2fp = max(i for i in range(ffma, len(data)) if src[i].startswith("F2FP") and int(data[i][idx["Instructions Executed"]]) > 100000)
sttm = first(lambda s: s.startswith("STTM.x16"), ffma)
pfull = first(lambda s: "SYNCS.ARRIVE" in s, sttm)
per_tile = int(data[ldtm][idx["Instructions Executed"]])
buckets = [("wait S (incl. first-tile TMA latency)", wait_s - 8, ldtm), ("tcgen05.ld S + s_free", ldtm, arrive + 1),
           ("mask (last tile only)", arrive + 1, fmn), ("row max + rescale decision", fmn, ffma),
           ("exp2 + row sum + pack", ffma, last_f2fp + 1), ("wait PV_(j-1) (+ O rescale)", last_f2fp + 1, sttm),
           ("tcgen05.st P + p_full", sttm, pfull + 1)]
tot = sum(S(r) for r in data)
print(f"{rows[0][1][:70]}: {tot} samples, {per_tile} warp-tiles")
for name, a, b in buckets:
    rs = data[a:b]
    s = sum(S(r) for r in rs)
    ex = sum(int(r[idx["Instructions Executed"]]) for r in rs) / per_tile
    agg = {k: sum(int(r[idx[k]]) for r in rs) for k in stalls}
    top = sorted(((v, k[6:]) for k, v in agg.items() if v), reverse=True)[:4]
    print(f"{name:40s} {s:6d} samples ({100 * s / tot:4.1f} %)  {ex:6.1f} instr per warp-tile  {top}")
agg = collections.defaultdict(collections.Counter)
for r, s in zip(data[ffma:last_f2fp + 1], src[ffma:last_f2fp + 1]):
    op = s.split()[1] if s.startswith("@") else s.split()[0]
    agg[op]["n"] += 1
    agg[op]["samples"] += S(r)
    for k in stalls:
        agg[op][k[6:]] += int(r[idx[k]])
print("inside the exp phase, by opcode:")
for op, c in sorted(agg.items(), key=lambda kv: -kv[1]["samples"])[:6]:
    print(f"  {op:24s} x{c['n']:4d} {c['samples']:6d} samples", [(k, v) for k, v in c.most_common(6) if k not in ("n", "samples")][:4])

"""Per-kernel count of the SASS mnemonics that tell a Blackwell-native kernel from a recompiled one
(B200_PROFILING.md): UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG/UTMAREDG = TMA load /
store / reduce, HMMA = mma.sync (legacy warp-level tensor path).  Reads the in-tree library with cuobjdump.

    python scripts/sass_summary.py > profiles/r02_sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vda_b200 import build

lib = build.build()
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
keys = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "HMMA", "MUFU.EX2", "LDGSTS"]
per = collections.OrderedDict()
cur = None
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = per.setdefault(re.sub(r"\(.*", "", name), collections.Counter())
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if not m:
        continue
    op = m.group(1)
    for k in keys:
        if op == k or op.startswith(k + "."):
            cur[k] += 1
    if op.startswith("UTCHMMA") and ".2CTA" in op:
        cur["UTCHMMA.2CTA"] += 0   # already counted by the prefix rule above when the suffix follows directly
print(f"library {os.path.relpath(lib)}  source digest {build.binary_digest()}")
print(f"{'kernel':70s} " + " ".join(f"{k:>12s}" for k in keys))
tot = collections.Counter()
for name, c in per.items():
    if sum(c.values()) == 0:
        continue
    tot.update(c)
    print(f"{name[:70]:70s} " + " ".join(f"{c[k]:12d}" for k in keys))
print(f"{'TOTAL':70s} " + " ".join(f"{tot[k]:12d}" for k in keys))

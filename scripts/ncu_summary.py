"""Key metrics of one .ncu-rep (ncu --set full) as JSON:
    ncu_summary.py rep out.json "command" "launch note" ['{"A": bytes, ...}' algorithmic bytes of the launch]
The summary is stamped with the digest of the GEMM sources in the tree (bench.py quotes a capture only while it
matches) and of the whole library."""
import csv
import io
import json
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

KEYS = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_active.avg", "smsp__cycles_active.avg"]
txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units, vals = rows[0], rows[1], rows[2]
out = {}
for k in KEYS:
    if k in hdr:
        i = hdr.index(k)
        v = vals[i].replace(",", "")
        try:
            v = float(v)
        except ValueError:
            pass
        out[k] = v if not units[i] or isinstance(v, str) else {"value": v, "unit": units[i]}
out["command"] = sys.argv[3] if len(sys.argv) > 3 else ""
out["launch"] = sys.argv[4] if len(sys.argv) > 4 else ""
if len(sys.argv) > 5:
    out["algorithmic_bytes"] = json.loads(sys.argv[5])
import bench
from vda_b200 import build

out["gemm_source_digest"] = bench.gemm_source_digest()
out["library_source_digest"] = build._digest()
json.dump(out, open(sys.argv[2], "w"), indent=1)
print(json.dumps(out, indent=1))

"""TMEM-read, MUFU.EX2 and F2FP rates per SM as the attention kernel's softmax warps see them (probe kernel in
scripts/probes/umma_probe.cu).  Prints bytes or instructions per clock per SM for 1, 2 and 4 warps per scheduler."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch

import vda_b200.lib as L
from probes.probes import load as load_probes

L.init()
PROBES = load_probes()
iters = 2000
names = {0: "tcgen05.ld 32x32b.x32 x4 (16 KB per warp-iteration)", 1: "ex2.approx.ftz.f32", 2: "cvt.rn.bf16x2.f32",
         3: "mix: 16 ex2 + 8 cvt + 8 fma.f32x2", 4: "max.f32 (3-input)", 5: "max.f32 (2-input)", 6: "fma.rn.f32x2",
         7: "add.f32x2", 8: "mix: 16 ex2 + 16 max3 (ex2 counted)"}
per_iter = {0: 16384, 1: 32 * 32, 2: 32 * 32, 3: 16 * 32, 4: 32 * 32, 5: 32 * 32, 6: 32 * 32, 7: 32 * 32, 8: 16 * 32}   # bytes / lane-ops (ex2 for the mix) per warp-iteration
for what in (tuple(int(a) for a in sys.argv[1:]) or (0, 1, 2, 3, 4, 5, 6, 7, 8)):
    for warps in (4, 8, 16):
        clk = torch.zeros(148 * warps, device="cuda", dtype=torch.int64)
        for _ in range(2):
            L.check(PROBES.vda_debug_pipe_rate(L.ptr(clk), what, 148, warps * 32, iters, L.stream_ptr()))
        torch.cuda.synchronize()
        c = clk.float().view(148, warps).max(dim=1).values.median().item()
        rate = per_iter[what] * warps * iters / c
        unit = "B/clk/SM" if what == 0 else "lane-ops/clk/SM"
        print(f"{names[what]:55s} warps/SM={warps:2d}: {c / iters:8.1f} clk per iteration, {rate:7.1f} {unit}", flush=True)

"""Peak issue rate of mma.sync.m16n8k16 bf16 (legacy warp-level tensor path) on this GPU."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch

import vda_b200.lib as L
from probes.probes import load as load_probes

L.init()
PROBES = load_probes()
out = torch.zeros(4, device="cuda")
iters = 20000
for warps in (4, 8, 12, 16):
    for ctas_per_sm in (1,):
        ctas = 148 * ctas_per_sm
        for _ in range(2):
            L.check(PROBES.vda_debug_mma_rate(L.ptr(out), ctas, warps * 32, iters, L.stream_ptr()))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        L.check(PROBES.vda_debug_mma_rate(L.ptr(out), ctas, warps * 32, iters, L.stream_ptr()))
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        n = ctas * warps * iters * 8
        print(f"warps/SM={warps}: {n * 4096 / ms / 1e9:.0f} TFLOP/s, {ms * 1e-3 * 1.8e9 / (iters * 8 * warps / 4):.2f} clk/mma/SMSP @1.8GHz", flush=True)

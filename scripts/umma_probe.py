"""Which (row0, sbo, base_offset) combinations does a tcgen05 SWIZZLE_128B K-major descriptor read correctly?"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch

import vda_b200.lib as L
from probes.probes import load as load_probes

L.init()
PROBES = load_probes()
g = torch.Generator(device="cuda").manual_seed(0)
G = torch.randn(320, 64, device="cuda", generator=g).bfloat16()
W = torch.randn(32, 64, device="cuda", generator=g).bfloat16()
D = torch.empty(128, 32, device="cuda")
for sbo in (1024, 2048, 1280):
    for row0 in (0, 1, 2, 8, 16, 17, 18, 33, 5):
        res = []
        for bo in sorted({0, row0 & 7}):
            D.fill_(float("nan"))
            L.check(PROBES.vda_debug_umma_shift(L.ptr(G), L.ptr(W), L.ptr(D), row0, sbo, bo, L.stream_ptr()))
            torch.cuda.synchronize()
            rows = torch.tensor([gidx * (sbo // 128) + row0 + i for gidx in range(16) for i in range(8)], device="cuda")
            ok = bool((rows < 320).all())
            want = G[rows.clamp(max=319)].float() @ W.float().T
            err = (D - want).abs().max().item()
            res.append(f"base_offset={bo}: max err {err:.3g}{'' if ok else ' (rows out of range)'}")
        print(f"sbo={sbo} row0={row0}: " + "; ".join(res), flush=True)

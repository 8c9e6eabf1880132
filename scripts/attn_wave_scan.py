"""Attention time vs CTA count around the 296-slot wave boundary (2 CTAs x 148 SMs): S=1370, heads=16,
frames T -> 11*16*T CTAs.  Shows how much the partial last wave of the 32-frame launch (5632 = 19*296 + 8) costs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vda_b200.lib as L
L.init()
heads, D, S = 16, 1024, 1370
spad = (S + 7) // 8 * 8
Tmax = 37
g = torch.Generator(device="cuda").manual_seed(0)
qk = torch.randn(Tmax * S, 2 * D, device="cuda", generator=g).bfloat16()
vt = torch.randn(Tmax * heads, 64, spad, device="cuda", generator=g).bfloat16()
ao = torch.empty(Tmax * S, D, device="cuda", dtype=torch.bfloat16)
for T in (26, 27, 28, 30, 31, 32, 33, 34, 35, 37):
    f = lambda: L.flash_attn(qk, vt, ao, T, S, heads, spad)
    for _ in range(3): f()
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): f()
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 10)
    ctas = 11 * heads * T
    print(f"T={T}: {best*1e3:7.1f} us ctas={ctas} waves={ctas/296:.2f} us/frame={best*1e3/T:.2f}", flush=True)

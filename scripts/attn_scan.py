"""Per-CTA fixed cost vs per-tile cost of the attention kernel: time(S) for a fixed CTA count per launch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vda_b200.lib as L
L.init()
heads, D = 16, 1024
def run(T, S):
    M = T * S; spad = (S + 7) // 8 * 8
    g = torch.Generator(device="cuda").manual_seed(0)
    qk = torch.randn(M, 2 * D, device="cuda", generator=g).bfloat16()
    vt = torch.randn(T * heads, 64, spad, device="cuda", generator=g).bfloat16()
    ao = torch.empty(M, D, device="cuda", dtype=torch.bfloat16)
    f = lambda: L.flash_attn(qk, vt, ao, T, S, heads, spad)
    for _ in range(3): f()
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): f()
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 10)
    qt = (S + 127) // 128
    ctas = qt * heads * T
    print(f"T={T} S={S}: {best*1e3:8.1f} us  ctas={ctas} tiles/cta={qt}  us per (cta-slot-tile) = {best*1e3*296/(ctas*qt):.3f}", flush=True)
for T, S in [(32, 1370), (32, 1408), (64, 640), (16, 2816), (128, 256), (8, 5632), (32, 128 * 11), (352, 128)]:
    run(T, S)

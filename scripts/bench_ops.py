"""Per-op timing of the vitl encoder block kernels with their real epilogues (CUDA events, 20 iters)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import vda_b200.lib as L

L.init()
T, S, D, heads = 32, 1370, 1024, 16
M = T * S
spad = (S + 7) // 8 * 8
dev = "cuda"
g = torch.Generator(device="cuda").manual_seed(0)
r = lambda *s, sc=1.0: (torch.randn(*s, device=dev, generator=g) * sc)
xn = r(M, D).bfloat16()
tok = r(M, D)
wqkv, bqkv = r(3 * D, D, sc=D ** -0.5).bfloat16(), r(3 * D)
wproj, bproj, g1 = r(D, D, sc=D ** -0.5).bfloat16(), r(D), r(D)
w1, b1 = r(4 * D, D, sc=D ** -0.5).bfloat16(), r(4 * D)
w2, b2 = r(D, 4 * D, sc=(4 * D) ** -0.5).bfloat16(), r(D)
qk = torch.empty(M, 2 * D, device=dev, dtype=torch.bfloat16)
vt = torch.zeros(T * heads, 64, spad, device=dev, dtype=torch.bfloat16)
ao = torch.empty(M, D, device=dev, dtype=torch.bfloat16)
h1 = torch.empty(M, 4 * D, device=dev, dtype=torch.bfloat16)
lnw, lnb = r(D), r(D)

ops = {
    "ln": (lambda: L.layernorm(tok, lnw, lnb, xn, 1e-6), 0),
    "qkv": (lambda: L.gemm(xn, wqkv, L.make_epilogue(bias=bqkv, out=qk, ldo=2 * D, store_mode=L.ST_QKV,
                                                     qkv=(vt, D, heads, S, spad))), 2 * M * 3 * D * D),
    "attn": (lambda: L.flash_attn(qk, vt, ao, T, S, heads, spad), 4 * T * heads * S * S * 64),
    "proj": (lambda: L.gemm(ao, wproj, L.make_epilogue(bias=bproj, scale=g1, res_f32=tok, out_f32=tok, ldo=D)),
             2 * M * D * D),
    "fc1": (lambda: L.gemm(xn, w1, L.make_epilogue(bias=b1, out=h1, ldo=4 * D, act=L.ACT_GELU)), 2 * M * 4 * D * D),
    "fc2": (lambda: L.gemm(h1, w2, L.make_epilogue(bias=b2, scale=g1, res_f32=tok, out_f32=tok, ldo=D)),
            2 * M * 4 * D * D),
    "fc1_bias": (lambda: L.gemm(xn, w1, L.make_epilogue(bias=b1, out=h1, ldo=4 * D)), 2 * M * 4 * D * D),
    "fc1_relu": (lambda: L.gemm(xn, w1, L.make_epilogue(bias=b1, out=h1, ldo=4 * D, act=L.ACT_RELU)), 2 * M * 4 * D * D),
    "fc1_plain": (lambda: L.gemm(xn, w1, L.make_epilogue(out=h1, ldo=4 * D)), 2 * M * 4 * D * D),
    "proj_plain": (lambda: L.gemm(ao, wproj, L.make_epilogue(out=xn, ldo=D)), 2 * M * D * D),
}
only = sys.argv[1:] or list(ops)
tot = 0.0
for name in only:
    fn, flops = ops[name]
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    tot += ms if "plain" not in name else 0
    print(f"{name:10s} {ms * 1e3:8.1f} us  {flops / ms / 1e9:7.0f} TFLOP/s")
print(f"sum of non-plain ops {tot:.3f} ms")

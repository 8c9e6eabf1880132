"""Per-op timing of the vitl encoder block kernels with their real epilogues (CUDA events, 20 iters)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import vda_b200.lib as L
from vda_b200 import packing as _P

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
aflags = torch.zeros(T * heads * ((S + 127) // 128), device=dev, dtype=torch.int32)
qkvs = (torch.randn(M, 9 * D, device=dev) * 0.5).bfloat16()   # FP32-accuracy mode: q | k | v as [hi | lo | hi]
ao3 = torch.empty(M, 3 * D, device=dev, dtype=torch.bfloat16)
h1 = torch.empty(M, 4 * D, device=dev, dtype=torch.bfloat16)
lnw, lnb = r(D), r(D)

# DPT-head 3x3 convolutions (refinenet level: 32 frames, 256 ch, 148x148 and 296x296 / 128->32 tail)
cx = r(T, 148, 148, 256).to(torch.bfloat16)
cw = r(256, 9 * 256).to(torch.bfloat16)
cb = r(256)
cy = torch.empty(T, 148, 148, 256, device=dev, dtype=torch.bfloat16)

# motion-module temporal attention (mm0: 37x37 tokens, C=1024; mm3: 74x74, C=256) and head upsamples
tq0 = r(T * 1369, 3 * 1024).to(torch.bfloat16)
to0 = torch.empty(T * 1369, 1024, device=dev, dtype=torch.bfloat16)
tq3 = r(T * 5476, 3 * 256).to(torch.bfloat16)
to3 = torch.empty(T * 5476, 256, device=dev, dtype=torch.bfloat16)
up_in = r(T, 148, 148, 256).to(torch.bfloat16)
up_out = torch.empty(T, 296, 296, 256, device=dev, dtype=torch.bfloat16)
up2_in = r(T, 296, 296, 128).to(torch.bfloat16)
up2_out = torch.empty(T, 518, 518, 128, device=dev, dtype=torch.bfloat16)

tw3, tb3 = _P.conv3x3(r(32, 128, 3, 3, sc=1152 ** -0.5), r(32), device=dev)
tw1 = r(32).abs()
tdepth = torch.empty(T, 518, 518, device=dev)

gn_x = r(T * 5476, 256).to(torch.bfloat16)
gn_o = torch.empty_like(gn_x)
gn_w, gn_b = r(256), r(256)
gn_ws = torch.empty(96 * T, device=dev, dtype=torch.float64)

oc_x = r(T, 296, 296, 256).to(torch.bfloat16)
oc_w, oc_b = _P.conv3x3(r(128, 256, 3, 3, sc=2304 ** -0.5), r(128), device=dev)
oc_y = torch.empty(T, 296, 296, 128, device=dev, dtype=torch.bfloat16)

# reassemble ConvTranspose k=s as GEMM + pixel-shuffle store (dpt.py:60-82): 37x37 tokens, 256 -> 256 x4x4 / 512 -> 512 x2x2
ct_a0 = r(T * 1369, 256).to(torch.bfloat16)
ct_w0, ct_b0 = r(16 * 256, 256, sc=1 / 16).to(torch.bfloat16), r(16 * 256)
ct_o0 = torch.empty(T, 148, 148, 256, device=dev, dtype=torch.bfloat16)
ct_a1 = r(T * 1369, 512).to(torch.bfloat16)
ct_w1, ct_b1 = r(4 * 512, 512, sc=1 / 22).to(torch.bfloat16), r(4 * 512)
ct_o1 = torch.empty(T, 74, 74, 512, device=dev, dtype=torch.bfloat16)

# motion module 3 (74x74 tokens, C=256): K = 256 GEMMs, epilogue-bound
m3_a = r(T * 5476, 256).to(torch.bfloat16)
m3_wqkv = r(768, 256, sc=1 / 16).to(torch.bfloat16)
m3_qkv = torch.empty(T * 5476, 768, device=dev, dtype=torch.bfloat16)
m3_wo, m3_bo = r(256, 256, sc=1 / 16).to(torch.bfloat16), r(256)
m3_h = r(T * 5476, 256)

m3_wg, m3_bg = r(2048, 256, sc=1 / 16).to(torch.bfloat16), r(2048)
m3_g = torch.empty(T * 5476, 1024, device=dev, dtype=torch.bfloat16)

cr1, cr2 = torch.randn_like(cy), torch.randn_like(cy)
cyr = torch.empty_like(cy)
m3_r1 = r(T * 5476, 256).to(torch.bfloat16)
m3_o = torch.empty_like(m3_r1)
pj_r1 = r(M, D).to(torch.bfloat16)
pj_o = torch.empty_like(pj_r1)

ops = {
    # bf16 residuals / ReLU'd second output (FeatureFusion / ResidualConvUnit epilogues, motion proj_out + residual)
    "conv148_r2": (lambda: L.conv3x3(cx, cw, L.make_epilogue(bias=cb, res1=cr1, res2=cr2, out=cy, out_relu=cyr, ldo=256), T, 148, 148, 256, 256),
                   2 * T * 148 * 148 * 256 * 9 * 256),
    "conv148_r1": (lambda: L.conv3x3(cx, cw, L.make_epilogue(bias=cb, res1=cr1, out=cy, ldo=256), T, 148, 148, 256, 256),
                   2 * T * 148 * 148 * 256 * 9 * 256),
    "conv148_orelu": (lambda: L.conv3x3(cx, cw, L.make_epilogue(bias=cb, out=cy, out_relu=cyr, ldo=256), T, 148, 148, 256, 256),
                      2 * T * 148 * 148 * 256 * 9 * 256),
    "mm3_res1": (lambda: L.gemm(m3_a, m3_wo, L.make_epilogue(bias=m3_bo, res1=m3_r1, out=m3_o, ldo=256)), 2 * T * 5476 * 256 * 256),
    "gemm1024_res1": (lambda: L.gemm(ao, wproj, L.make_epilogue(bias=bproj, res1=pj_r1, out=pj_o, ldo=D)), 2 * M * D * D),
    "mm3_geglu": (lambda: L.gemm(m3_a, m3_wg, L.make_epilogue(bias=m3_bg, out=m3_g, ldo=1024, act=L.ACT_GEGLU)), 2 * T * 5476 * 2048 * 256),
    "mm3_qkv": (lambda: L.gemm(m3_a, m3_wqkv, L.make_epilogue(out=m3_qkv, ldo=768)), 2 * T * 5476 * 768 * 256),
    "mm3_out": (lambda: L.gemm(m3_a, m3_wo, L.make_epilogue(bias=m3_bo, res_f32=m3_h, out_f32=m3_h, ldo=256)), 2 * T * 5476 * 256 * 256),
    "ct0": (lambda: L.gemm(ct_a0, ct_w0, L.make_epilogue(bias=ct_b0, out=ct_o0, ldo=256, store_mode=L.ST_CONVT, ct=(4, 256, 37, 37))),
            2 * T * 1369 * 4096 * 256),
    "ct1": (lambda: L.gemm(ct_a1, ct_w1, L.make_epilogue(bias=ct_b1, out=ct_o1, ldo=512, store_mode=L.ST_CONVT, ct=(2, 512, 37, 37))),
            2 * T * 1369 * 2048 * 512),
    "oc1": (lambda: L.conv3x3(oc_x, oc_w, L.make_epilogue(bias=oc_b, out=oc_y, ldo=128), T, 296, 296, 256, 128),
            2 * T * 296 * 296 * 128 * 2304),
    "oc1_1cta": (lambda: L.conv3x3(oc_x, oc_w, L.make_epilogue(bias=oc_b, out=oc_y, ldo=128, cta_pair=1), T, 296, 296, 256, 128),
                 2 * T * 296 * 296 * 128 * 2304),
    "tailconv": (lambda: L.conv3x3(up2_out, tw3, L.make_epilogue(bias=tb3, out_f32=tdepth, w2=tw1, b2=0.1,
                                                                  store_mode=L.ST_HEADDOT), T, 518, 518, 128, 32),
                 2 * T * 518 * 518 * 32 * 1152),
    "tail_fused": (lambda: L.depth_tail(up2_in, tw3, tb3, tw1, 0.1, tdepth, T, 296, 296, 518, 518, 128), 2 * T * 518 * 518 * 32 * 1152),
    "gn74": (lambda: L.groupnorm_nhwc(gn_x, gn_w, gn_b, gn_o, gn_ws, T, 5476, 256, 1e-6), 0),
    "tail_unfused": (lambda: (L.bilinear_nhwc(up2_in, up2_out, T, 296, 296, 518, 518, 128),
                              L.conv3x3(up2_out, tw3, L.make_epilogue(bias=tb3, out_f32=tdepth, w2=tw1, b2=0.1,
                                                                      store_mode=L.ST_HEADDOT), T, 518, 518, 128, 32)),
                     2 * T * 518 * 518 * 32 * 1152),
    "tattn0": (lambda: L.temporal_attn(tq0, 3072, tq0[:, 1024:], 3072, tq0[:, 2048:], 3072, to0, 1024, 1369, 8, 128, 32, 32),
               4 * 1369 * 8 * 32 * 32 * 128),
    "tattn3": (lambda: L.temporal_attn(tq3, 768, tq3[:, 256:], 768, tq3[:, 512:], 768, to3, 256, 5476, 8, 32, 32, 32),
               4 * 5476 * 8 * 32 * 32 * 32),
    "up296": (lambda: L.bilinear_nhwc(up_in, up_out, T, 148, 148, 296, 296, 256), 0),
    "up518": (lambda: L.bilinear_nhwc(up2_in, up2_out, T, 296, 296, 518, 518, 128), 0),
    "conv148": (lambda: L.conv3x3(cx, cw, L.make_epilogue(bias=cb, out=cy, ldo=256, act=L.ACT_RELU), T, 148, 148, 256, 256),
                2 * T * 148 * 148 * 256 * 9 * 256),
    "ln": (lambda: L.layernorm(tok, lnw, lnb, xn, 1e-6), 0),
    "qkv": (lambda: L.gemm(xn, wqkv, L.make_epilogue(bias=bqkv, out=qk, ldo=2 * D, store_mode=L.ST_QKV,
                                                     qkv=(vt, D, heads, S, spad))), 2 * M * 3 * D * D),
    "attn_f32": (lambda: L.attn_f32(qkvs, ao3, D, T, S, heads, in_split=3 * D), 3 * 4 * T * heads * S * S * 64),
    "attn": (lambda: L.flash_attn(qk, vt, ao, T, S, heads, spad, aflags), 4 * T * heads * S * S * 64),
    "attn_exact": (lambda: L.flash_attn(qk, vt, ao, T, S, heads, spad), 4 * T * heads * S * S * 64),
    "proj": (lambda: L.gemm(ao, wproj, L.make_epilogue(bias=bproj, scale=g1, res_f32=tok, out_f32=tok, ldo=D)),
             2 * M * D * D),
    "fc1": (lambda: L.gemm(xn, w1, L.make_epilogue(bias=b1, out=h1, ldo=4 * D, act=L.ACT_GELU)), 2 * M * 4 * D * D),
    "fc2": (lambda: L.gemm(h1, w2, L.make_epilogue(bias=b2, scale=g1, res_f32=tok, out_f32=tok, ldo=D)),
            2 * M * 4 * D * D),
    "fc1_bias": (lambda: L.gemm(xn, w1, L.make_epilogue(bias=b1, out=h1, ldo=4 * D)), 2 * M * 4 * D * D),
    "fc1_relu": (lambda: L.gemm(xn, w1, L.make_epilogue(bias=b1, out=h1, ldo=4 * D, act=L.ACT_RELU)), 2 * M * 4 * D * D),
    "proj_wf32": (lambda: L.gemm(ao, wproj, L.make_epilogue(bias=bproj, scale=g1, out_f32=tok, ldo=D)), 2 * M * D * D),
    "proj_rres": (lambda: L.gemm(ao, wproj, L.make_epilogue(bias=bproj, scale=g1, res_f32=tok, out=xn, ldo=D)), 2 * M * D * D),
    "proj_bias": (lambda: L.gemm(ao, wproj, L.make_epilogue(bias=bproj, scale=g1, out=xn, ldo=D)), 2 * M * D * D),
    "fc1_plain": (lambda: L.gemm(xn, w1, L.make_epilogue(out=h1, ldo=4 * D)), 2 * M * 4 * D * D),
    "proj_plain": (lambda: L.gemm(ao, wproj, L.make_epilogue(out=xn, ldo=D)), 2 * M * D * D),
}
# usage: bench_ops.py [op ...] [--env NAME=v1,v2,...]  -> every op is timed under each value of the
# environment knob, interleaved round-robin over 5 rounds; the best round of each is reported.
args = sys.argv[1:]
env_name, env_vals = None, [None]
if "--env" in args:
    i = args.index("--env")
    env_name, vals = args[i + 1].split("=")
    env_vals = vals.split(",")
    del args[i:i + 2]
only = args or list(ops)


def timed(fn, iters=15):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


best = {}
for rnd in range(5):
    for name in only:
        for val in env_vals:
            if env_name:
                os.environ[env_name] = val
            fn = ops[name][0]
            fn()
            ms = timed(fn)
            best[(name, val)] = min(best.get((name, val), 1e9), ms)
tot = {v: 0.0 for v in env_vals}
for name in only:
    for val in env_vals:
        ms, flops = best[(name, val)], ops[name][1]
        tot[val] += ms if "plain" not in name and "bias" not in name and "relu" not in name and "f32" not in name else 0
        tag = f" [{env_name}={val}]" if env_name else ""
        print(f"{name:10s} {ms * 1e3:8.1f} us  {flops / ms / 1e9:7.0f} TFLOP/s{tag}")
for val in env_vals:
    print(f"sum of block ops {tot[val]:.3f} ms" + (f" [{env_name}={val}]" if env_name else ""))

"""A few SMALL launches of every hand-written kernel family, for compute-sanitizer (memcheck / racecheck / synccheck
each in its own gpurun call):  compute-sanitizer --tool <tool> python scripts/sanitize_small.py
Results are checked against torch so a run that is "clean" but wrong still fails."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

import vda_b200.lib as L

L.init()
g = torch.Generator().manual_seed(0)
r = lambda *s, sc=1.0: (torch.randn(*s, generator=g) * sc).cuda()


def close(got, want, tol=2e-2, what=""):
    err = (got.float() - want.float()).abs().max().item()
    ref = want.float().abs().max().item()
    print(f"{what}: max err {err:.3e} (ref max {ref:.3e})", flush=True)
    assert err <= tol * max(ref, 1e-6), what


# ---- GEMM: 1-CTA and CTA-pair kernels, bf16 box store, fp32 reduce-add (with tail split-K shape), GEGLU, QKV store
for pair in (1, 2):
    M, N, K = 300, 256, 192
    a, w, b = r(M, K).bfloat16(), r(N, K, sc=K ** -0.5).bfloat16(), r(N)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    L.gemm(a, w, L.make_epilogue(bias=b, out=out, ldo=N, act=L.ACT_GELU, cta_pair=pair))
    close(out, F.gelu(a.float() @ w.float().T + b), what=f"gemm gelu pair={pair}")
    res = r(M, N)
    want = res + (a.float() @ w.float().T + b) * 0.5
    L.gemm(a, w, L.make_epilogue(bias=b, scale=torch.full((N,), 0.5, device="cuda"), res_f32=res, out_f32=res, ldo=N, cta_pair=pair))
    close(res, want, what=f"gemm reduce-add pair={pair}")
M, N, K = 128 * 2 * 80, 256, 512     # 80 pair tiles on 74 pairs: the partial last wave is cut into K ranges
a, w, b = r(M, K).bfloat16(), r(N, K, sc=K ** -0.5).bfloat16(), r(N)
res = r(M, N)
want = res + a.float() @ w.float().T + b
L.gemm(a, w, L.make_epilogue(bias=b, res_f32=res, out_f32=res, ldo=N))
close(res, want, what="gemm tail split-K")
frames, S, heads = 2, 197, 2
D = heads * 64
spad = (S + 7) // 8 * 8
a, w, b = r(frames * S, D).bfloat16(), r(3 * D, D, sc=D ** -0.5).bfloat16(), r(3 * D)
qk = torch.empty(frames * S, 2 * D, device="cuda", dtype=torch.bfloat16)
vt = torch.zeros(frames * heads, 64, spad, device="cuda", dtype=torch.bfloat16)
L.gemm(a, w, L.make_epilogue(bias=b, out=qk, ldo=2 * D, store_mode=L.ST_QKV, qkv=(vt, D, heads, S, spad)))
qkv = a.float() @ w.float().T + b
close(qk, qkv[:, :2 * D], what="qkv store")
# ---- spatial attention on the GEMM's outputs
ao = torch.empty(frames * S, D, device="cuda", dtype=torch.bfloat16)
L.flash_attn(qk, vt, ao, frames, S, heads, spad)
f = qkv.view(frames, S, 3, heads, 64).permute(2, 0, 3, 1, 4)
close(ao, F.scaled_dot_product_attention(f[0], f[1], f[2]).permute(0, 2, 1, 3).reshape(frames * S, D), what="attention")
# ---- 3x3 convs: tap-by-tap pair kernel, pair halo mode, 1-CTA halo mode with the fused 32-wide dot
for (n, H, W, C, Cout, kw, tag) in ((1, 40, 40, 64, 256, dict(cta_pair=2), "conv pair"),
                                    (2, 120, 120, 64, 128, dict(cta_pair=2, block_n=128), "conv pair halo"),
                                    (1, 37, 21, 128, 64, dict(), "conv 1-cta halo")):
    x, wt, b = r(n, C, H, W).bfloat16(), r(Cout, C, 3, 3, sc=(9 * C) ** -0.5).bfloat16(), r(Cout)
    out = torch.empty(n, H, W, Cout, device="cuda", dtype=torch.bfloat16)
    L.conv3x3(x.permute(0, 2, 3, 1).contiguous(), wt.permute(0, 2, 3, 1).reshape(Cout, 9 * C).contiguous(),
              L.make_epilogue(bias=b, out=out, ldo=Cout, **kw), n, H, W, C, Cout)
    close(out, F.conv2d(x.float(), wt.float(), b, padding=1).permute(0, 2, 3, 1), what=tag)
n, H, W, C = 1, 42, 56, 64
x, wt, b, w2 = r(n, C, H, W).bfloat16(), r(32, C, 3, 3, sc=(9 * C) ** -0.5).bfloat16(), r(32), r(32)
out = torch.empty(n, H, W, device="cuda")
L.conv3x3(x.permute(0, 2, 3, 1).contiguous(), wt.permute(0, 2, 3, 1).reshape(32, 9 * C).contiguous(),
          L.make_epilogue(bias=b, out_f32=out, w2=w2, b2=0.3, store_mode=L.ST_HEADDOT), n, H, W, C, 32)
mid = F.conv2d(x.float(), wt.float(), b, padding=1).relu()
close(out, (torch.einsum("nchw,c->nhw", mid, w2) + 0.3).relu(), what="conv headdot")
# ---- LayerNorm, temporal attention, bilinear, alignment kernels
x = r(500, 384)
gm, bt = r(384), r(384)
o = torch.empty(500, 384, device="cuda", dtype=torch.bfloat16)
L.layernorm(x, gm, bt, o, 1e-6)
close(o, F.layer_norm(x, (384,), gm, bt, 1e-6), what="layernorm")
T, hw, Cc = 32, 21, 256
q3 = r(T * hw, 3 * Cc).bfloat16()
to = torch.empty(T * hw, Cc, device="cuda", dtype=torch.bfloat16)
L.temporal_attn(q3, 3 * Cc, q3[:, Cc:], 3 * Cc, q3[:, 2 * Cc:], 3 * Cc, to, Cc, hw, 8, Cc // 8, T, T)
qq = q3.float().view(T, hw, 3, 8, Cc // 8).permute(2, 1, 3, 0, 4)
close(to.view(T, hw, 8, Cc // 8), F.scaled_dot_product_attention(qq[0], qq[1], qq[2]).permute(2, 0, 1, 3), what="temporal attention")
xb = r(2, 20, 30, 64).bfloat16()
ob = torch.empty(2, 35, 53, 64, device="cuda", dtype=torch.bfloat16)
L.bilinear_nhwc(xb, ob, 2, 20, 30, 35, 53, 64)
close(ob, F.interpolate(xb.float().permute(0, 3, 1, 2), size=(35, 53), mode="bilinear", align_corners=True).permute(0, 2, 3, 1),
      what="bilinear")
p_, t_ = r(2, 33, 47).abs() + 0.5, r(2, 33, 47).abs() + 0.5
sums = torch.zeros(5, device="cuda", dtype=torch.float64)
ss = torch.zeros(2, device="cuda")
L.lsq_accum(p_, t_, p_.numel(), sums)
L.lsq_solve(sums, ss, reset=True)
dst = torch.empty_like(p_)
L.affine_clamp(p_, dst, p_.numel(), ss)
A = torch.stack([p_.flatten(), torch.ones_like(p_.flatten())], 1).double()
sol = torch.linalg.lstsq(A, t_.flatten().double()[:, None]).solution.flatten()
close(ss, sol.float(), 1e-4, what="lsq")
torch.cuda.synchronize()
print("sanitize_small: all kernels ran and matched")

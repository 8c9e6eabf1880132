"""FP32-accuracy path (infer_video_depth(..., fp32=True), video_depth.py:108).

Tolerance (BASELINE.json north_star): "the FP32 path must agree within rtol 1e-3" with the reference
PyTorch implementation.  The check is np.allclose-style: |got - want| <= RTOL*|want| + ATOL with
ATOL = 1e-5 * max|want|."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import make_golden as G
from oracle import refload
from oracle import vda_oracle as O

pytestmark = pytest.mark.gpu
RTOL = 1e-3


@pytest.fixture(scope="module")
def L():
    import vda_b200.lib as lib

    lib.init()
    return lib


@pytest.fixture(autouse=True)
def _exact_fp32_oracle():
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    yield


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(*shape, device="cuda", generator=g) * scale


def split3(x):
    """f32 [rows, C] -> bf16 [rows, 3C] = [hi | lo | hi]."""
    hi = x.bfloat16()
    lo = (x - hi.float()).bfloat16()
    return torch.cat([hi, lo, hi], 1).contiguous()


def unsplit(t):
    c = t.shape[1] // 3
    assert torch.equal(t[:, :c], t[:, 2 * c:])
    return t[:, :c].float() + t[:, c:2 * c].float()


def split_close(got_split, want, extra=0.0):
    """A split tensor represents x to 2^-18 |x|; allow 3x that (plus `extra` for the op's own rounding)."""
    got = unsplit(got_split)
    tol = 1.2e-5 * want.abs().max().item() + extra
    err = (got - want).abs().max().item()
    assert err <= tol, f"max err {err:.3e} > {tol:.3e}"


def check_close(got, want, rtol=RTOL, what=""):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    atol = 1e-5 * np.abs(want).max()
    err = np.abs(got - want) - (rtol * np.abs(want) + atol)
    worst = float((np.abs(got - want) / (np.abs(want) + atol)).max())
    print(f"{what}: worst rel err {worst:.2e}")
    assert (err <= 0).all(), f"{what}: {int((err > 0).sum())} of {err.size} elements outside rtol {rtol}"


def test_split_gemm_is_fp32_accurate(L):
    from vda_b200 import packing as P

    M, N, K = 300, 96, 520
    a, w, b = _rand(M, K, seed=1), _rand(N, K, seed=2, scale=K ** -0.5), _rand(N, seed=3)
    res = _rand(M, N, seed=4)
    with P.split_mode(True):
        wk, bk = P.linear(w, b)
    out = torch.empty(M, 3 * N, device="cuda", dtype=torch.bfloat16)
    outr = torch.empty_like(out)
    o32 = torch.empty(M, N, device="cuda")
    L.gemm(split3(a), wk, L.make_epilogue(bias=bk, res1=split3(res), out=out, out_relu=outr, ldo=3 * N, split=N))
    L.gemm(split3(a), wk, L.make_epilogue(bias=bk, res_f32=res, out_f32=o32, ldo=N))
    want = (a.double() @ w.double().T + b.double() + res.double()).float()
    # a product of split operands carries ~2^-16 relative error of the |a||w| sum
    tol = 3e-5 * (a.abs() @ w.abs().T).max().item()
    assert (unsplit(out) - want).abs().max().item() < tol
    assert (unsplit(outr) - want.relu()).abs().max().item() < tol
    assert (o32 - want).abs().max().item() < tol


def test_split_conv3x3_and_geglu(L):
    from vda_b200 import packing as P

    n, H, W, C, Co = 2, 21, 19, 64, 64
    x, w, b = _rand(n, C, H, W, seed=1), _rand(Co, C, 3, 3, seed=2, scale=(9 * C) ** -0.5), _rand(Co, seed=3)
    with P.split_mode(True):
        wk, bk = P.conv3x3(w, b)
    xs = split3(x.permute(0, 2, 3, 1).reshape(-1, C))
    out = torch.empty(n * H * W, 3 * Co, device="cuda", dtype=torch.bfloat16)
    L.conv3x3(xs, wk, L.make_epilogue(bias=bk, out=out, ldo=3 * Co, split=Co, act=L.ACT_RELU), n, H, W, 3 * C, Co)
    want = F.conv2d(x.double(), w.double(), b.double(), padding=1).relu().permute(0, 2, 3, 1).reshape(-1, Co).float()
    split_close(out, want, 2e-5)
    # GEGLU (motion_module/attention.py:382-384)
    M, Cc = 200, 64
    a, wg, bg = _rand(M, Cc, seed=5), _rand(8 * Cc, Cc, seed=6, scale=Cc ** -0.5), _rand(8 * Cc, seed=7)
    with P.split_mode(True):
        wgk, bgk = P.geglu(wg, bg, "cuda")
    g = torch.empty(M, 3 * 4 * Cc, device="cuda", dtype=torch.bfloat16)
    L.gemm(split3(a), wgk, L.make_epilogue(bias=bgk, out=g, ldo=12 * Cc, split=4 * Cc, act=L.ACT_GEGLU))
    y = a.double() @ wg.double().T + bg.double()
    want = (y[:, :4 * Cc] * F.gelu(y[:, 4 * Cc:])).float()
    split_close(g, want, 3e-5)


def test_split_elementwise_kernels(L):
    rows, C = 77, 192
    x = _rand(rows, C, seed=1) * 3 + 0.5
    g, b = _rand(C, seed=2), _rand(C, seed=3)
    out = torch.empty(rows, 3 * C, device="cuda", dtype=torch.bfloat16)
    L.layernorm(x, g, b, out, 1e-6, split=1)
    want = F.layer_norm(x.double(), (C,), g.double(), b.double(), 1e-6).float()
    split_close(out, want)
    # GroupNorm on a split NHWC map
    n, hw, Cg = 3, 50, 128
    xm = _rand(n * hw, Cg, seed=4) * 2 + 1
    gg, gb = _rand(Cg, seed=5), _rand(Cg, seed=6)
    og = torch.empty(n * hw, 3 * Cg, device="cuda", dtype=torch.bfloat16)
    stats = torch.empty(96 * n, device="cuda", dtype=torch.float64)
    L.groupnorm_nhwc(split3(xm), gg, gb, og, stats, n, hw, Cg, 1e-6, ldx=3 * Cg, split=Cg)
    want = F.group_norm(xm.double().view(n, hw, Cg).permute(0, 2, 1), 32, gg.double(), gb.double(), 1e-6)
    want = want.permute(0, 2, 1).reshape(-1, Cg).float()
    split_close(og, want, 1e-5)
    # bilinear (align_corners=True)
    n, h, w, H, W, Cb = 2, 9, 11, 18, 23, 64
    xb = _rand(n, Cb, h, w, seed=7)
    ob = torch.empty(n * H * W, 3 * Cb, device="cuda", dtype=torch.bfloat16)
    L.bilinear_nhwc(split3(xb.permute(0, 2, 3, 1).reshape(-1, Cb)), ob, n, h, w, H, W, Cb, ldx=3 * Cb, ldo=3 * Cb,
                    split_in=Cb, split_out=Cb)
    want = F.interpolate(xb.double(), size=(H, W), mode="bilinear", align_corners=True).permute(0, 2, 3, 1).reshape(-1, Cb)
    split_close(ob, want.float())
    # patch gather
    xi = _rand(2, 3, 28, 42, seed=8)
    a = torch.empty(2 * 2 * 3, 3 * 592, device="cuda", dtype=torch.bfloat16)
    L.patch_im2col(xi, a, 2, 28, 42, 592, split=1)
    want = F.unfold(xi, 14, stride=14).permute(0, 2, 1).reshape(-1, 588)
    split_close(a[:, [c + 592 * p for p in range(3) for c in range(588)]], want)
    assert unsplit(a)[:, 588:].abs().max().item() == 0


@pytest.mark.parametrize("frames,S,heads", [(2, 1370, 6), (3, 131, 2), (1, 64, 1), (1, 7, 1)])
def test_attn_f32(L, frames, S, heads):
    """q/k/v arrive as the split tensor the projection GEMM writes."""
    D = heads * 64
    qkv = _rand(frames * S, 3 * D, seed=1)
    out = torch.full((frames * S, 3 * D), float("nan"), device="cuda", dtype=torch.bfloat16)
    L.attn_f32(split3(qkv), out, D, frames, S, heads, in_split=3 * D)
    f = qkv.double().view(frames, S, 3, heads, 64).permute(2, 0, 3, 1, 4)
    want = F.scaled_dot_product_attention(f[0], f[1], f[2]).permute(0, 2, 1, 3).reshape(frames * S, D).float()
    split_close(out, want, 1e-5)


@pytest.mark.parametrize("T,hw,C,Tq", [(32, 40, 1024, 32), (32, 37, 256, 32), (4, 35, 64, 4), (7, 20, 384, 7)])
def test_temporal_attn_f32(L, T, hw, C, Tq):
    heads, hd = 8, C // 8
    qkv = _rand(T * hw, 3 * C, seed=1)
    out = torch.full((Tq * hw, 3 * C), float("nan"), device="cuda", dtype=torch.bfloat16)
    L.temporal_attn_f32(qkv, C, out, C, hw, heads, Tq, T)
    f = qkv.double().view(T, hw, 3, heads, hd).permute(2, 1, 3, 0, 4)
    want = F.scaled_dot_product_attention(f[0][:, :, :Tq], f[1], f[2]).permute(2, 0, 1, 3).reshape(Tq * hw, C).float()
    split_close(out, want, 1e-5)


def build(arch, seed=0, metric=False):
    from video_depth_anything.video_depth import VideoDepthAnything

    cfg = O.ARCH[arch]
    m = VideoDepthAnything(encoder=arch, features=cfg["features"], out_channels=cfg["out_channels"], metric=metric)
    m.load_state_dict(refload.seeded_state_dict(arch, seed), strict=True)
    return m.to("cuda").eval()


@pytest.mark.parametrize("name", sorted(G.FORWARD_CASES))
def test_fp32_forward_vs_reference_golden(name):
    """Golden depth maps written by the unmodified reference (fp32, CPU): oracle/make_golden.py."""
    arch, T, H, W = G.FORWARD_CASES[name]
    model = build(arch)
    x = G.window_input(T, H, W).cuda()
    with torch.no_grad():
        got = model(x, fp32=True).cpu().numpy()
    want = np.load(os.path.join(G.GOLDEN_DIR, name + ".npz"))["depth"]
    assert got.shape == want.shape and (got >= 0).all()
    check_close(got, want, what=name)


@pytest.mark.parametrize("arch,T,H,W", [("vits", 32, 518, 518), ("vitl", 32, 518, 518), ("vitb", 4, 518, 924)])
def test_fp32_full_size_window_vs_oracle_on_gpu(arch, T, H, W):
    model = build(arch)
    sd_gpu = {k: v.cuda() for k, v in refload.seeded_state_dict(arch, 0).items()}
    x = G.window_input(T, H, W, seed=3).cuda()
    with torch.no_grad():
        got = model(x, fp32=True)[0]
        want = O.forward(sd_gpu, x, arch)[0]
    check_close(got.cpu().numpy(), want.cpu().numpy(), what=f"{arch} {T}x{H}x{W}")


def test_fp32_video_driver_vs_reference_golden():
    v = G.VIDEO_CASE
    frames = G.video_frames(v["n"], v["h"], v["w"])
    gold = np.load(os.path.join(G.GOLDEN_DIR, "video_vits_46x44x60.npz"))
    model = build(v["arch"])
    got, _ = model.infer_video_depth(frames, 24, input_size=v["input_size"], device="cuda", fp32=True)
    check_close(got, gold["offline"], what="driver fp32")


def test_fp32_streaming_vs_reference_golden():
    """infer_video_depth_one(..., fp32=True): fp32 K/V caches + fp32 cached-attention kernel, frame by frame
    against the reference's streaming outputs (46 frames: past the cache's steady state)."""
    from video_depth_anything.video_depth_stream import VideoDepthAnything as Stream

    v = G.VIDEO_CASE
    frames = G.video_frames(v["n"], v["h"], v["w"])
    gold = np.load(os.path.join(G.GOLDEN_DIR, "video_vits_46x44x60.npz"))["stream"]
    cfg = O.ARCH[v["arch"]]
    m = Stream(encoder=v["arch"], features=cfg["features"], out_channels=cfg["out_channels"])
    m.load_state_dict(refload.seeded_state_dict(v["arch"], 0), strict=True)
    m = m.to("cuda").eval()
    got = np.stack([m.infer_video_depth_one(frames[i], input_size=v["input_size"], device="cuda", fp32=True)
                    for i in range(v["n"])])
    check_close(got, gold, what="stream fp32")
    with pytest.raises(RuntimeError):      # the cache holds fp32 projections: no switching mid-stream
        m.infer_video_depth_one(frames[0], input_size=v["input_size"], device="cuda", fp32=False)

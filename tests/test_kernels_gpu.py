"""Attention / normalisation / resampling / alignment kernels vs plain torch fp32 references."""
import numpy as np
import os

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    import vda_b200.lib as lib

    lib.init()
    return lib


def _rand(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).cuda()


def _close(got, want, tol=2e-2):
    got, want = got.float(), want.float()
    assert torch.isfinite(got).all()
    err = (got - want).abs().max().item()
    ref = want.abs().max().item()
    assert err <= tol * max(ref, 1e-6), f"max abs err {err} vs ref max {ref}"


# the optimistic-softmax variant exists in experiment builds only (VDA_NVCC_EXTRA=-DVDA_ATTN_EXPERIMENTS)
_EXPERIMENTS = "VDA_ATTN_EXPERIMENTS" in os.environ.get("VDA_NVCC_EXTRA", "")


def _attn_flags(frames, S, heads):
    return torch.zeros(frames * heads * ((S + 127) // 128), device="cuda", dtype=torch.int32)


@pytest.mark.parametrize("optimistic", [False, True])
@pytest.mark.parametrize("frames,S,heads", [(1, 128, 1), (2, 197, 2), (3, 1370, 6), (1, 2443, 3), (2, 64, 1), (32, 1370, 16), (1, 7, 2)])
def test_flash_attention(L, frames, S, heads, optimistic, monkeypatch):
    """optimistic: offset fixed after the first KV tile + fix-up pass (flags stay zero on well-behaved scores);
    opt-in with VDA_ATTN_OPT=1 (slower than the exact kernel on B200, kept as a measured variant)."""
    if optimistic and not _EXPERIMENTS:
        pytest.skip("optimistic softmax: experiment builds only")
    monkeypatch.setenv("VDA_ATTN_OPT", "1" if optimistic else "0")
    D = heads * 64
    spad = (S + 7) // 8 * 8
    q = _rand(frames, S, heads, 64, scale=1.5, seed=1).bfloat16()
    k = _rand(frames, S, heads, 64, scale=1.5, seed=2).bfloat16()
    v = _rand(frames, S, heads, 64, seed=3).bfloat16()
    qk = torch.cat([q.reshape(frames * S, D), k.reshape(frames * S, D)], dim=1).contiguous()
    vt = torch.zeros(frames * heads, 64, spad, device="cuda", dtype=torch.bfloat16)
    vt[:, :, :S] = v.permute(0, 2, 3, 1).reshape(frames * heads, 64, S)
    out = torch.full((frames * S, D), float("nan"), device="cuda", dtype=torch.bfloat16)
    flags = _attn_flags(frames, S, heads) if optimistic else None
    L.flash_attn(qk, vt, out, frames, S, heads, spad, flags)
    torch.cuda.synchronize()
    want = F.scaled_dot_product_attention(q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2))
    _close(out.view(frames, S, heads, 64), want.transpose(1, 2), 2e-2)
    if optimistic:
        assert int(flags.sum()) == 0


@pytest.mark.parametrize("boost", [6.0, 60.0, 400.0])
def test_flash_attention_optimistic_fixup(L, boost, monkeypatch):
    """Scores far above the first KV tile's maximum: a moderate jump stays inside the optimistic kernel's safe range,
    a huge one overflows its fixed offset, raises the tile's flag and is recomputed exactly by the fix-up pass.  Either
    way the result is the exact softmax and the flags are zero again afterwards."""
    if not _EXPERIMENTS:
        pytest.skip("optimistic softmax: experiment builds only")
    monkeypatch.setenv("VDA_ATTN_OPT", "1")
    frames, S, heads = 2, 700, 2
    D = heads * 64
    spad = (S + 7) // 8 * 8
    q = _rand(frames, S, heads, 64, scale=1.0, seed=1).bfloat16()
    k = _rand(frames, S, heads, 64, scale=1.0, seed=2)
    k[:, 300:340] *= boost            # keys of the third tile dominate (for the rows whose q.k is positive)
    k[:, 600:] *= 0.5 * boost
    k = k.bfloat16()
    v = _rand(frames, S, heads, 64, seed=3).bfloat16()
    qk = torch.cat([q.reshape(frames * S, D), k.reshape(frames * S, D)], dim=1).contiguous()
    vt = torch.zeros(frames * heads, 64, spad, device="cuda", dtype=torch.bfloat16)
    vt[:, :, :S] = v.permute(0, 2, 3, 1).reshape(frames * heads, 64, S)
    out = torch.full((frames * S, D), float("nan"), device="cuda", dtype=torch.bfloat16)
    flags = _attn_flags(frames, S, heads)
    L.flash_attn(qk, vt, out, frames, S, heads, spad, flags)
    torch.cuda.synchronize()
    want = F.scaled_dot_product_attention(q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2))
    _close(out.view(frames, S, heads, 64), want.transpose(1, 2), 2e-2)
    assert int(flags.sum()) == 0, "the fix-up pass must consume every flag"
    ref = torch.full_like(out, float("nan"))
    L.flash_attn(qk, vt, ref, frames, S, heads, spad)          # exact kernel alone
    assert (out.float() - ref.float()).abs().max().item() <= 2e-2 * want.abs().max().item()


def test_flash_attention_growing_maxima(L):
    """Key magnitudes grow along the sequence, so the running row maximum jumps by far more than the
    lazy-rescale threshold (2^8) at several KV tiles: exercises the in-TMEM rescale of O."""
    frames, S, heads = 2, 1000, 2
    D = heads * 64
    spad = (S + 7) // 8 * 8
    q = _rand(frames, S, heads, 64, scale=2.0, seed=1).bfloat16()
    ramp = torch.linspace(0.2, 6.0, S, device="cuda").view(1, S, 1, 1)
    k = (_rand(frames, S, heads, 64, seed=2) * ramp).bfloat16()
    v = _rand(frames, S, heads, 64, seed=3).bfloat16()
    qk = torch.cat([q.reshape(frames * S, D), k.reshape(frames * S, D)], dim=1).contiguous()
    vt = torch.zeros(frames * heads, 64, spad, device="cuda", dtype=torch.bfloat16)
    vt[:, :, :S] = v.permute(0, 2, 3, 1).reshape(frames * heads, 64, S)
    out = torch.full((frames * S, D), float("nan"), device="cuda", dtype=torch.bfloat16)
    L.flash_attn(qk, vt, out, frames, S, heads, spad)
    torch.cuda.synchronize()
    want = F.scaled_dot_product_attention(q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2))
    _close(out.view(frames, S, heads, 64), want.transpose(1, 2), 2e-2)


@pytest.mark.parametrize("rows,C,f32", [(100, 384, True), (2740, 1024, True), (77, 256, False), (50, 64, False), (33, 768, True), (10, 192, False)])
def test_layernorm(L, rows, C, f32):
    x = _rand(rows, C, scale=2.0, seed=1) + 0.5
    if not f32:
        x = x.bfloat16()
    g, b = _rand(C, seed=2), _rand(C, seed=3)
    out = torch.empty(rows, C, device="cuda", dtype=torch.bfloat16)
    L.layernorm(x, g, b, out, 1e-6)
    _close(out, F.layer_norm(x.float(), (C,), g, b, 1e-6), 1e-2)


def test_layernorm_skip_cls_and_pe(L):
    frames, S, C = 4, 11, 384
    x = _rand(frames * S, C, seed=1)
    g, b = _rand(C, seed=2), _rand(C, seed=3)
    want = F.layer_norm(x, (C,), g, b, 1e-6).view(frames, S, C)[:, 1:].reshape(-1, C)
    out = torch.empty(frames * (S - 1), C, device="cuda", dtype=torch.bfloat16)
    L.layernorm(x, g, b, out, 1e-6, tokens_per_frame=S, skip=1)
    _close(out, want, 1e-2)
    T, hw = 4, 11
    pe = _rand(32, C, seed=4)
    out = torch.empty(T * hw, C, device="cuda", dtype=torch.bfloat16)
    nope = torch.empty_like(out)
    L.layernorm(x, g, b, out, 1e-5, pe=pe, out_nope=nope, pe_rows_per_t=hw, pe_T=32)
    base = F.layer_norm(x, (C,), g, b, 1e-5)
    _close(nope, base, 1e-2)
    _close(out, (base.view(T, hw, C) + pe[:T, None]).view(-1, C), 1e-2)


@pytest.mark.parametrize("n,h,w,C", [(3, 19, 19, 1024), (2, 37, 37, 256), (4, 5, 7, 192), (2, 10, 10, 64), (2, 37, 37, 384), (3, 19, 19, 768),
                                     (1, 74, 74, 128), (32, 37, 37, 1024)])
def test_groupnorm_nhwc(L, n, h, w, C):
    x = (_rand(n, C, h, w, scale=2.0, seed=1) + 1.0).bfloat16()
    g, b = _rand(C, seed=2), _rand(C, seed=3)
    want = F.group_norm(x.float(), 32, g, b, 1e-6).permute(0, 2, 3, 1)
    xn = x.permute(0, 2, 3, 1).contiguous()
    out = torch.empty(n, h, w, C, device="cuda", dtype=torch.bfloat16)
    ws = torch.empty(96 * n, device="cuda", dtype=torch.float64)
    L.groupnorm_nhwc(xn, g, b, out, ws, n, h * w, C, 1e-6)
    _close(out, want, 1e-2)


def test_patch_embed_path(L):
    frames, H, W, D = 2, 42, 70, 384
    x = _rand(frames, 3, H, W, seed=1)
    wt = _rand(D, 3, 14, 14, scale=588 ** -0.5, seed=2)
    b = _rand(D, seed=3)
    cls, pos = _rand(D, seed=4), _rand(1 + 15, D, seed=5)
    hw = 15
    a = torch.empty(frames * hw, 592, device="cuda", dtype=torch.bfloat16)
    L.patch_im2col(x, a, frames, H, W, 592)
    wk = torch.zeros(D, 592, device="cuda", dtype=torch.bfloat16)
    wk[:, :588] = wt.reshape(D, 588)
    patch = torch.empty(frames * hw, D, device="cuda")
    L.gemm(a, wk, L.make_epilogue(bias=b, out_f32=patch, ldo=D), K=588)
    tokens = torch.empty(frames, hw + 1, D, device="cuda")
    L.add_cls_pos(patch, cls, pos, tokens, frames, hw, D)
    want = F.conv2d(x, wt, b, stride=14).flatten(2).transpose(1, 2)
    want = torch.cat([cls.expand(frames, 1, D), want], dim=1) + pos
    _close(tokens, want, 1e-2)


@pytest.mark.parametrize("h,w,H,W,C", [(19, 19, 37, 37, 64), (37, 66, 74, 132, 128), (296, 296, 518, 518, 8), (5, 5, 5, 5, 16)])
def test_bilinear_nhwc(L, h, w, H, W, C):
    n = 2
    x = _rand(n, C, h, w, seed=1).bfloat16()
    want = F.interpolate(x.float(), size=(H, W), mode="bilinear", align_corners=True).permute(0, 2, 3, 1)
    out = torch.empty(n, H, W, C, device="cuda", dtype=torch.bfloat16)
    L.bilinear_nhwc(x.permute(0, 2, 3, 1).contiguous(), out, n, h, w, H, W, C)
    _close(out, want, 1e-2)


def test_bilinear_f32(L):
    x = _rand(3, 1, 70, 98, seed=1)
    out = torch.empty(3, 123, 77, device="cuda")
    L.bilinear_f32(x, out, 3, 70, 98, 123, 77, relu=True)
    want = F.interpolate(x, size=(123, 77), mode="bilinear", align_corners=True).relu()[:, 0]
    assert (out - want).abs().max().item() < 1e-5


def test_im2col_s2_conv(L):
    n, h, w, C, Co = 2, 37, 37, 64, 64
    x = _rand(n, C, h, w, seed=1).bfloat16()
    wt = _rand(Co, C, 3, 3, scale=(9 * C) ** -0.5, seed=2).bfloat16()
    b = _rand(Co, seed=3)
    want = F.conv2d(x.float(), wt.float(), b, stride=2, padding=1).permute(0, 2, 3, 1)
    ho = wo = 19
    a = torch.empty(n * ho * wo, 9 * C, device="cuda", dtype=torch.bfloat16)
    L.im2col_s2(x.permute(0, 2, 3, 1).contiguous(), a, n, h, w, C)
    out = torch.empty(n, ho, wo, Co, device="cuda", dtype=torch.bfloat16)
    L.gemm(a, wt.permute(0, 2, 3, 1).reshape(Co, 9 * C).contiguous(), L.make_epilogue(bias=b, out=out, ldo=Co))
    _close(out, want)


@pytest.mark.parametrize("T,hw,C,Tq", [(32, 50, 1024, 32), (32, 37, 256, 32), (4, 35, 192, 4), (32, 12, 384, 1), (32, 9, 64, 32), (7, 20, 128, 7)])
def test_temporal_attention(L, T, hw, C, Tq):
    heads, hd = 8, C // 8
    qkv = _rand(T * hw, 3 * C, seed=1).bfloat16()
    q = qkv[(T - Tq) * hw:, :C]
    out = torch.full((Tq * hw, C), float("nan"), device="cuda", dtype=torch.bfloat16)
    L.temporal_attn(q, 3 * C, qkv[:, C:], 3 * C, qkv[:, 2 * C:], 3 * C, out, C, hw, heads, hd, Tq, T)
    f = qkv.float().view(T, hw, 3, heads, hd).permute(2, 1, 3, 0, 4)       # [3, hw, heads, T, hd]
    want = F.scaled_dot_product_attention(f[0][:, :, T - Tq:], f[1], f[2])   # [hw, heads, Tq, hd]
    _close(out.view(Tq, hw, heads, hd), want.permute(2, 0, 1, 3), 2e-2)


def test_lsq_and_blend(L):
    from oracle import vda_oracle as O

    rng = np.random.default_rng(0)
    h, w = 60, 80
    pred = rng.uniform(0.5, 8, (2, h, w)).astype(np.float32)
    target = (pred * 1.7 - 0.4 + rng.normal(0, 0.05, pred.shape)).astype(np.float32)
    s_ref, t_ref = O.scale_and_shift(pred.reshape(2 * h, w), target.reshape(2 * h, w))
    sums = torch.zeros(5, device="cuda", dtype=torch.float64)
    ss = torch.zeros(2, device="cuda")
    p, t = torch.from_numpy(pred).cuda(), torch.from_numpy(target).cuda()
    L.lsq_accum(p[0], t[0], h * w, sums)
    L.lsq_accum(p[1], t[1], h * w, sums)
    L.lsq_solve(sums, ss)
    assert abs(ss[0].item() - s_ref) < 1e-4 * abs(s_ref) and abs(ss[1].item() - t_ref) < 1e-3
    L.lsq_solve(sums, ss, metric=True)
    assert ss.tolist() == [1.0, 0.0]
    ss = torch.tensor([1.3, -2.0], device="cuda")
    pre = rng.uniform(0, 5, (8, h, w)).astype(np.float32)
    post = rng.uniform(0, 5, (8, h, w)).astype(np.float32)
    fixed = [O._affine_clamp(post[i].copy(), np.float32(1.3), np.float32(-2.0)) for i in range(8)]
    wts = [0.0] + [i * (1.0 / 7) for i in range(1, 7)] + [1.0]
    want = np.stack([pre[i] * (1 - wts[i]) + fixed[i] * wts[i] for i in range(8)])
    pre_d, post_d = torch.from_numpy(pre).cuda(), torch.from_numpy(post).cuda()
    L.affine_clamp_blend(pre_d, post_d, h * w, 8, ss)
    np.testing.assert_allclose(pre_d.cpu().numpy(), want, rtol=1e-6, atol=1e-6)
    dst = torch.empty_like(post_d)
    L.affine_clamp(post_d, dst, post_d.numel(), ss)
    np.testing.assert_allclose(dst.cpu().numpy(), np.stack(fixed), rtol=1e-6, atol=1e-6)


def test_attention_throughput_report(L, capsys):
    frames, S, heads = 32, 1370, 16
    D, spad = heads * 64, 1376
    qk = _rand(frames * S, 2 * D, seed=1).bfloat16()
    vt = _rand(frames * heads, 64, spad, seed=2).bfloat16()
    out = torch.empty(frames * S, D, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        L.flash_attn(qk, vt, out, frames, S, heads, spad)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(10):
        L.flash_attn(qk, vt, out, frames, S, heads, spad)
    t1.record()
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / 10
    fl = 4 * frames * heads * S * S * 64
    with capsys.disabled():
        print(f"\n[attn] vitl block: {ms:.3f} ms = {fl / ms / 1e9:.0f} TFLOP/s")


@pytest.mark.parametrize("h0,w0,size", [(90, 120, 112), (720, 1280, 518), (518, 518, 518), (300, 200, 518)])
def test_preprocess_u8_matches_cv2_path(L, h0, w0, size):
    """GPU pre-processing against the reference's own host path (cv2 INTER_CUBIC + float64 normalise)."""
    from vda_b200.preprocess import net_input_size, preprocess_frames, preprocess_frames_cv2

    rng = np.random.default_rng(0)
    frames = rng.integers(0, 256, (3, h0, w0, 3), dtype=np.uint8)
    nh, nw = net_input_size(h0, w0, size)
    got = preprocess_frames(frames, [2, 0], nh, nw, "cuda")
    want = preprocess_frames_cv2(frames, [2, 0], nh, nw, "cuda")
    assert got.shape == want.shape == (2, 3, nh, nw)
    err = (got - want).abs().max().item()
    print(f"preprocess {h0}x{w0}->{nh}x{nw}: max abs diff vs cv2 {err:.2e}")
    assert err < 1e-3  # cv2's float cubic path (IPP / SIMD) differs in the last bits of the tap weights: <= 1e-4 of the input range


@pytest.mark.parametrize("n,h,w,H,W,C", [(2, 40, 40, 70, 70, 128), (3, 24, 40, 42, 70, 64), (1, 296, 296, 518, 518, 128),
                                         (2, 296, 528, 518, 924, 64), (1, 8, 16, 14, 28, 128)])
def test_depth_tail_fused(L, n, h, w, H, W, C):
    """Fused upsample + 3x3(->32) + ReLU + 1x1 + ReLU (tcgen05, csrc/depth_tail.cu) against the unfused kernel pair
    (same bf16 interpolation to the bit, so only the fp32 summation order differs) and against torch."""
    from vda_b200 import packing as P

    assert L.depth_tail_supported(h, w, H, W, C)
    x = _rand(n, C, h, w, seed=1).bfloat16()
    w3 = _rand(32, C, 3, 3, scale=(9 * C) ** -0.5, seed=2)
    b3, w1 = _rand(32, seed=3), _rand(32, seed=4).abs()
    b1 = 0.05
    xn = x.permute(0, 2, 3, 1).contiguous()
    wk, bk = P.conv3x3(w3, b3, device="cuda")
    numel, guard = n * H * W, 4096
    buf = torch.full((numel + 2 * guard,), 7.0, device="cuda")
    out = buf[guard:guard + numel].view(n, H, W)
    out.fill_(float("nan"))
    L.depth_tail(xn, wk, bk, w1, b1, out, n, h, w, H, W, C)
    up = torch.empty(n, H, W, C, device="cuda", dtype=torch.bfloat16)
    L.bilinear_nhwc(xn, up, n, h, w, H, W, C)
    ref2 = torch.empty(n, H, W, device="cuda")
    L.conv3x3(up, wk, L.make_epilogue(bias=bk, out_f32=ref2, w2=w1, b2=b1, store_mode=L.ST_HEADDOT), n, H, W, C, 32)
    torch.cuda.synchronize()
    assert not torch.isnan(out).any(), "some output pixels were never written"
    assert (buf[:guard] == 7.0).all() and (buf[guard + numel:] == 7.0).all(), "store outside the output tensor"
    upt = F.interpolate(x.float(), size=(H, W), mode="bilinear", align_corners=True).bfloat16().float()
    y = F.conv2d(upt, w3.bfloat16().float(), b3, padding=1).relu()
    want = (y * w1.view(1, 32, 1, 1)).sum(1).add(b1).relu()
    scale = want.abs().max().item()
    assert (out - ref2).abs().max().item() < 2e-3 * scale      # same operands, different summation order
    assert (out - want).abs().max().item() < 2e-2 * scale      # torch rounds the upsample differently in the last bf16 bit


def test_depth_colorize_bit_exact(L):
    """dc_utils.py:72-86: clip-global min/max, three float32 roundings, truncating uint8 cast -- byte for byte --
    then the colour table look-up.  Includes negative values (min/max are taken over the raw floats)."""
    from oracle import vda_oracle as O
    from utils.dc_utils import depth_to_vis, inferno_lut

    rng = np.random.default_rng(0)
    for d in (rng.uniform(0.2, 7.5, (3, 37, 53)).astype(np.float32),
              rng.normal(0.0, 30.0, (2, 518, 518)).astype(np.float32),
              (rng.integers(0, 1 << 16, (4, 64, 64)) / 257.0).astype(np.float32)):
        q = O.depth_to_uint8(d)
        got_grey = depth_to_vis(d, grayscale=True)
        assert got_grey.dtype == np.uint8 and got_grey.shape == q.shape
        np.testing.assert_array_equal(got_grey, q)
        got = depth_to_vis(d)
        assert got.shape == d.shape + (3,)
        np.testing.assert_array_equal(got, inferno_lut()[q])


def test_depth_normalize_bit_exact(L):
    """external_engine/engine.py:1099-1141 per frame: near/far remap, far <= near, invert_depth, constant frame."""
    import os

    from oracle import make_golden as G
    from oracle import vda_oracle as O

    gold = np.load(os.path.join(G.GOLDEN_DIR, "util_cases.npz"))
    depth = G.util_inputs()[5]
    big = np.random.default_rng(3).uniform(0, 50, (5, 270, 480)).astype(np.float32)
    big[3] = -1.5
    for i, (near, far, inv) in enumerate(G.NORMALIZE_CASES):
        got = L.depth_normalize(torch.from_numpy(depth).cuda(), near, far, inv).cpu().numpy()
        np.testing.assert_array_equal(got, gold[f"norm{i}"])          # the unmodified reference's output
        got = L.depth_normalize(torch.from_numpy(big).cuda(), near, far, inv).cpu().numpy()
        want = np.stack([O.normalize_depth(big[f], near, far, inv) for f in range(big.shape[0])])
        np.testing.assert_array_equal(got, want)


def test_util_shims_match_reference_golden(L):
    """utils.util.compute_scale_and_shift (masked, all-ones, scale_only) and get_interpolate_frames through the
    GPU kernels vs the unmodified reference's outputs (tests/golden/util_cases.npz)."""
    import os

    from oracle import make_golden as G
    from utils.util import compute_scale_and_shift, get_interpolate_frames

    gold = np.load(os.path.join(G.GOLDEN_DIR, "util_cases.npz"))
    pred, target, mask, pre, post, _ = G.util_inputs()
    np.testing.assert_allclose(np.array(compute_scale_and_shift(pred, target, mask), np.float64), gold["full"], rtol=2e-5)
    np.testing.assert_allclose(np.array(compute_scale_and_shift(pred, target, np.ones_like(mask)), np.float64),
                               gold["full_ones"], rtol=2e-5)
    np.testing.assert_allclose(np.array(compute_scale_and_shift(pred, target), np.float64), gold["full_ones"], rtol=2e-5)
    np.testing.assert_allclose(np.array(compute_scale_and_shift(pred, target, mask, scale_only=True), np.float64),
                               gold["scale_only"], rtol=2e-5)
    np.testing.assert_array_equal(np.stack(get_interpolate_frames(pre, post)), gold["blend"])
    # lsq_solve(reset=True) leaves zeroed sums behind
    sums = torch.ones(5, device="cuda", dtype=torch.float64)
    ss = torch.zeros(2, device="cuda")
    L.lsq_solve(sums, ss, reset=True)
    assert sums.abs().sum().item() == 0.0


@pytest.mark.parametrize("hw,C,Tk,f32", [(1369, 1024, 32, False), (361, 1024, 32, False), (5476, 256, 32, False), (35, 64, 32, False),
                                        (35, 192, 7, False), (130, 384, 1, False), (77, 768, 20, False), (50, 128, 32, True),
                                        (361, 1024, 32, True), (35, 48 * 8, 13, True)])
def test_temporal_attn_stream(L, hw, C, Tk, f32):
    """Streaming temporal attention (video_depth_stream.py:134-153, motion_module.py:252-266): the new frame attends over
    Tk list positions whose PE-free K / V projections sit in slot-addressed caches; checked against torch, incl. the
    cache insert of the new frame."""
    heads, hd, cap = 8, C // 8, 44
    dt = torch.float32 if f32 else torch.bfloat16
    qkv = _rand(hw, 3 * C, seed=1).to(dt)
    kc, vc = _rand(hw, cap, C, seed=2).to(dt), _rand(hw, cap, C, seed=3).to(dt)   # [hw, cap, C]
    kc0, vc0 = kc.clone(), vc.clone()
    pe_q, pe_k, pe_v = _rand(32, C, scale=0.3, seed=4), _rand(32, C, scale=0.3, seed=5), _rand(32, C, scale=0.3, seed=6)
    g = torch.Generator().manual_seed(7)
    slots = torch.randperm(cap, generator=g)[:32].to(torch.int32).cuda()
    q, kn, vn = qkv[:, :C].float(), qkv[:, C:2 * C].float(), qkv[:, 2 * C:].float()
    sl = slots[:Tk].long()
    K = kc0[:, sl].float().transpose(0, 1).clone()
    V = vc0[:, sl].float().transpose(0, 1).clone()
    K[Tk - 1], V[Tk - 1] = kn, vn
    qq = (q + pe_q[Tk - 1]).view(hw, heads, hd)
    KK = (K + pe_k[:Tk, None]).view(Tk, hw, heads, hd)
    VV = (V + pe_v[:Tk, None]).view(Tk, hw, heads, hd)
    s = torch.einsum("nhd,tnhd->nht", qq, KK) * hd ** -0.5
    want = torch.einsum("nht,tnhd->nhd", s.softmax(-1), VV).reshape(hw, C)
    if f32:
        out = torch.full((hw, 3 * C), float("nan"), device="cuda", dtype=torch.bfloat16)
        L.temporal_attn_stream_f32(qkv, qkv[:, C:], qkv[:, 2 * C:], 3 * C, kc, vc, slots, pe_q, pe_k, pe_v, out, C, hw, heads, hd, Tk)
        got = out[:, :C].float() + out[:, C:2 * C].float()
        assert torch.equal(out[:, :C], out[:, 2 * C:])
        _close(got, want, 1e-4)
    else:
        out = torch.full((hw, C), float("nan"), device="cuda", dtype=torch.bfloat16)
        L.temporal_attn_stream(qkv, qkv[:, C:], qkv[:, 2 * C:], 3 * C, kc, vc, slots, pe_q, pe_k, pe_v, out, C, hw, heads, hd, Tk)
        _close(out, want, 1e-2)
    torch.cuda.synchronize()
    new = int(slots[Tk - 1])
    assert torch.equal(kc[:, new], qkv[:, C:2 * C]) and torch.equal(vc[:, new], qkv[:, 2 * C:])
    keep = torch.ones(cap, dtype=torch.bool)
    keep[new] = False
    assert torch.equal(kc[:, keep], kc0[:, keep]) and torch.equal(vc[:, keep], vc0[:, keep])

"""tcgen05 GEMM / implicit-GEMM kernel vs a plain torch fp32 reference of the same op."""
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
    err = (got - want).abs().max().item()
    ref = want.abs().max().item()
    assert err <= tol * max(ref, 1e-6), f"max abs err {err} vs ref max {ref}"


@pytest.mark.parametrize("M,N,K,bn", [
    (128, 256, 64, 256), (256, 256, 128, 0), (300, 384, 592, 0), (1000, 64, 1024, 64),
    (1370, 1024, 1024, 128), (129, 32, 64, 32), (4096, 3072, 1024, 256), (777, 48, 256, 0),
])
def test_gemm_plain_bias(L, M, N, K, bn):
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    b = _rand(N, seed=3)
    out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    L.gemm(a, w, L.make_epilogue(bias=b, out=out, ldo=N, block_n=bn, cta_pair=1))
    torch.cuda.synchronize()
    _close(out, a.float() @ w.float().T + b)


@pytest.mark.parametrize("M,N,K,bn", [
    (128, 256, 64, 256), (256, 256, 128, 256), (300, 384, 592, 128), (1000, 64, 1024, 64), (129, 64, 64, 64),
    (1370, 1024, 1024, 128), (4096 + 128, 3072, 1024, 256), (777, 48, 256, 64), (43840, 1024, 1024, 0),
])
def test_gemm_cta_pair_kernel(L, M, N, K, bn):
    """tcgen05.mma.cta_group::2 kernel forced (odd tile counts, ragged M/N, partial pair)."""
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    b, g = _rand(N, seed=3), _rand(N, seed=4)
    out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    L.gemm(a, w, L.make_epilogue(bias=b, scale=g, out=out, ldo=N, block_n=bn, cta_pair=2))
    torch.cuda.synchronize()
    _close(out, (a.float() @ w.float().T + b) * g)


@pytest.mark.parametrize("n,H,W,C,Cout", [(2, 19, 19, 64, 64), (3, 37, 37, 128, 256), (1, 5, 200, 192, 128)])
def test_conv3x3_cta_pair_kernel(L, n, H, W, C, Cout):
    x = _rand(n, C, H, W, seed=1).bfloat16()
    wt = _rand(Cout, C, 3, 3, scale=(9 * C) ** -0.5, seed=2).bfloat16()
    b = _rand(Cout, seed=3)
    want = F.conv2d(x.float(), wt.float(), b, padding=1).permute(0, 2, 3, 1)
    xn = x.permute(0, 2, 3, 1).contiguous()
    wk = wt.permute(0, 2, 3, 1).reshape(Cout, 9 * C).contiguous()
    out = torch.full((n, H, W, Cout), float("nan"), device="cuda", dtype=torch.bfloat16)
    L.conv3x3(xn, wk, L.make_epilogue(bias=b, out=out, ldo=Cout, cta_pair=2), n, H, W, C, Cout)
    _close(out, want)


def test_gemm_gelu_scale_residual_inplace(L):
    M, N, K = 1370 * 2, 384, 1536
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    b, g = _rand(N, seed=3), _rand(N, seed=4)
    x = _rand(M, N, seed=5)
    want = x + g * (a.float() @ w.float().T + b)
    L.gemm(a, w, L.make_epilogue(bias=b, scale=g, res_f32=x, out_f32=x, ldo=N))
    torch.cuda.synchronize()
    _close(x, want, 1e-3)
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    L.gemm(a, w, L.make_epilogue(bias=b, out=out, ldo=N, act=L.ACT_GELU))
    _close(out, F.gelu(a.float() @ w.float().T + b))


@pytest.mark.parametrize("M,N,K", [
    (256 * 84, 256, 1024),      # 84 pair tiles on 74 SM pairs: 10 tail tiles, each cut into 4 K ranges
    (256 * 96 + 77, 256, 4096),  # 97 tiles (ragged last pair): 23 tail tiles x 3 ranges
    (43840, 1024, 1024),        # encoder proj: 688 tiles, 22 tail tiles x 3 ranges of 5/5/6 K blocks
])
def test_gemm_residual_reduce_add_tail_split_k(L, M, N, K):
    """In-place fp32 residual update x += g*(a@w.T + b) on the CTA-pair kernel where the last, partial wave
    of tiles is split along K (GemmParams::ksplit): partial sums meet in x through the TMA reduce-add."""
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    b, g = _rand(N, seed=3), _rand(N, seed=4)
    x = _rand(M, N, seed=5)
    want = x + g * (a.float() @ w.float().T + b)
    L.gemm(a, w, L.make_epilogue(bias=b, scale=g, res_f32=x, out_f32=x, ldo=N))
    torch.cuda.synchronize()
    _close(x, want, 1e-3)


def test_gemm_bf16_residuals_and_relu_copy(L):
    M, N, K = 999, 256, 256
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    b = _rand(N, seed=3)
    r1, r2 = _rand(M, N, seed=4).bfloat16(), _rand(M, N, seed=5).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    out_relu = torch.empty_like(out)
    L.gemm(a, w, L.make_epilogue(bias=b, res1=r1, res2=r2, out=out, out_relu=out_relu, ldo=N))
    want = a.float() @ w.float().T + b + r1.float() + r2.float()
    _close(out, want)
    _close(out_relu, want.relu())


@pytest.mark.parametrize("M,C", [(1500, 256), (256 * 80 + 5, 256), (5000, 64), (300, 1024)])
def test_gemm_geglu(L, M, C):
    """value * gelu(gate) epilogue: box path on the 1-CTA (BN=128) and CTA-pair (BN=256) kernels, generic path for BN=64."""
    a = _rand(M, C, seed=1).bfloat16()
    w = _rand(8 * C, C, scale=C ** -0.5, seed=2).bfloat16()
    b = _rand(8 * C, seed=3)
    want = a.float() @ w.float().T + b
    want = want[:, :4 * C] * F.gelu(want[:, 4 * C:])
    # interleave value/gate rows in groups of 16 (what the weight repack does)
    idx = torch.arange(8 * C, device="cuda").view(2, 4 * C // 16, 16).permute(1, 0, 2).reshape(-1)
    out = torch.full((M, 4 * C), float("nan"), device="cuda", dtype=torch.bfloat16)
    L.gemm(a, w[idx].contiguous(), L.make_epilogue(bias=b[idx].contiguous(), out=out, ldo=4 * C, act=L.ACT_GEGLU))
    _close(out, want)


@pytest.mark.parametrize("M,N,K", [(1370, 1024, 4096), (1370, 1024, 1024), (361, 1024, 4096), (130, 256, 512), (5476, 256, 1024)])
def test_gemm_small_m_split_k(L, M, N, K):
    """Small-M reduce-add GEMMs (streaming proj / fc2): every pair tile is cut into K ranges that meet in the fp32
    residual (api.cu: small_m_ksplit); bias must be added exactly once, LayerScale applies to every partial."""
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    b, g = _rand(N, seed=3), _rand(N, seed=4)
    res = _rand(M, N, seed=5)
    want = res + (a.float() @ w.float().T + b) * g
    L.gemm(a, w, L.make_epilogue(bias=b, scale=g, res_f32=res, out_f32=res, ldo=N))
    torch.cuda.synchronize()
    _close(res, want, 2e-3)


@pytest.mark.parametrize("halo2", ["1", "2", "0"])
@pytest.mark.parametrize("n,H,W,C,Cout,bn", [(2, 120, 120, 128, 128, 128), (1, 296, 296, 256, 128, 128), (3, 21, 37, 64, 128, 128),
                                             (2, 50, 70, 192, 256, 256), (1, 148, 148, 256, 256, 256)])
def test_conv3x3_cta_pair_halo(L, n, H, W, C, Cout, bn, halo2, monkeypatch):
    """Halo mode of the CTA-pair kernel (one 10x18 patch per channel block + W ring): BN = 128 by default,
    BN = 256 with VDA_CONV_HALO2=2; ragged tile edges, odd tile counts; 0 = the tap-by-tap path as comparand."""
    monkeypatch.setenv("VDA_CONV_HALO2", halo2)
    x = _rand(n, C, H, W, seed=1).bfloat16()
    wt = _rand(Cout, C, 3, 3, scale=(9 * C) ** -0.5, seed=2).bfloat16()
    b = _rand(Cout, seed=3)
    want = F.conv2d(x.float(), wt.float(), b, padding=1).relu().permute(0, 2, 3, 1)
    xn = x.permute(0, 2, 3, 1).contiguous()
    wk = wt.permute(0, 2, 3, 1).reshape(Cout, 9 * C).contiguous()
    numel, guard = n * H * W * Cout, 4096
    buf = torch.full((numel + 2 * guard,), 7.0, device="cuda", dtype=torch.bfloat16)
    out = buf[guard:guard + numel].view(n, H, W, Cout)
    out.fill_(float("nan"))
    L.conv3x3(xn, wk, L.make_epilogue(bias=b, out=out, ldo=Cout, act=L.ACT_RELU, cta_pair=2, block_n=bn), n, H, W, C, Cout)
    torch.cuda.synchronize()
    _close(out, want)
    assert (buf[:guard] == 7.0).all() and (buf[guard + numel:] == 7.0).all(), "store outside the output tensor"


@pytest.mark.parametrize("s,cin,cout,h,w", [(4, 64, 48, 5, 7), (2, 128, 96, 9, 4), (4, 256, 256, 37, 37), (2, 128, 64, 9, 4),
                                             (2, 512, 512, 37, 66), (4, 64, 32, 19, 3)])
def test_gemm_convtranspose_store(L, s, cin, cout, h, w):
    n = 3
    x = _rand(n, cin, h, w, seed=1).bfloat16()
    wt = _rand(cin, cout, s, s, scale=cin ** -0.5, seed=2).bfloat16()
    b = _rand(cout, seed=3)
    want = F.conv_transpose2d(x.float(), wt.float(), b, stride=s)  # [n,cout,h*s,w*s]
    cp = (cout + 15) // 16 * 16
    wk = torch.zeros(s, s, cp, cin, device="cuda", dtype=torch.bfloat16)
    wk[:, :, :cout] = wt.permute(2, 3, 1, 0)
    bk = torch.zeros(s, s, cp, device="cuda")
    bk[:, :, :cout] = b
    a = x.permute(0, 2, 3, 1).reshape(n * h * w, cin).contiguous()
    # guard bands of a sentinel value on both sides of the output: the scattered (pixel-shuffle) stores must not leave it
    numel, guard = n * h * s * w * s * cp, 4096
    buf = torch.full((numel + 2 * guard,), 7.0, device="cuda", dtype=torch.bfloat16)
    out = buf[guard:guard + numel].view(n, h * s, w * s, cp)
    out.fill_(float("nan"))
    L.gemm(a, wk.view(-1, cin), L.make_epilogue(bias=bk.view(-1), out=out, ldo=cp, store_mode=L.ST_CONVT,
                                                 ct=(s, cp, h, w)))
    torch.cuda.synchronize()
    assert not torch.isnan(out).any(), "some output pixels were never written"
    _close(out[..., :cout].permute(0, 3, 1, 2), want)
    assert out[..., cout:].abs().max().item() == 0 if cp > cout else True
    assert (buf[:guard] == 7.0).all() and (buf[guard + numel:] == 7.0).all(), "store outside the output tensor"


def test_gemm_qkv_store(L):
    frames, S, heads = 3, 197, 6
    D = heads * 64
    spad = (S + 7) // 8 * 8
    a = _rand(frames * S, D, seed=1).bfloat16()
    w = _rand(3 * D, D, scale=D ** -0.5, seed=2).bfloat16()
    b = _rand(3 * D, seed=3)
    qk = torch.empty(frames * S, 2 * D, device="cuda", dtype=torch.bfloat16)
    vt = torch.zeros(frames * heads, 64, spad, device="cuda", dtype=torch.bfloat16)
    L.gemm(a, w, L.make_epilogue(bias=b, out=qk, ldo=2 * D, store_mode=L.ST_QKV, qkv=(vt, D, heads, S, spad)))
    want = a.float() @ w.float().T + b
    _close(qk, want[:, :2 * D])
    v = want[:, 2 * D:].view(frames, S, heads, 64).permute(0, 2, 3, 1).reshape(frames * heads, 64, S)
    _close(vt[:, :, :S], v)
    assert vt[:, :, S:].abs().max().item() == 0


@pytest.mark.parametrize("n,H,W,C,Cout", [(2, 19, 19, 64, 64), (1, 37, 37, 128, 256), (3, 8, 50, 64, 32),
                                          (2, 74, 74, 256, 256), (1, 5, 200, 192, 48)])
def test_conv3x3(L, n, H, W, C, Cout):
    x = _rand(n, C, H, W, seed=1).bfloat16()
    wt = _rand(Cout, C, 3, 3, scale=(9 * C) ** -0.5, seed=2).bfloat16()
    b = _rand(Cout, seed=3)
    r1 = _rand(n, H, W, Cout, seed=4).bfloat16()
    want = F.conv2d(x.float(), wt.float(), b, padding=1).permute(0, 2, 3, 1) + r1.float()
    xn = x.permute(0, 2, 3, 1).contiguous()
    wk = wt.permute(0, 2, 3, 1).reshape(Cout, 9 * C).contiguous()
    out = torch.full((n, H, W, Cout), float("nan"), device="cuda", dtype=torch.bfloat16)
    out_relu = torch.empty_like(out)
    L.conv3x3(xn, wk, L.make_epilogue(bias=b, res1=r1, out=out, out_relu=out_relu, ldo=Cout), n, H, W, C, Cout)
    _close(out, want)
    _close(out_relu, want.relu())


def test_conv3x3_headdot(L):
    n, H, W, C = 2, 42, 56, 64
    x = _rand(n, C, H, W, seed=1).bfloat16()
    wt = _rand(32, C, 3, 3, scale=(9 * C) ** -0.5, seed=2).bfloat16()
    b, w2 = _rand(32, seed=3), _rand(32, seed=4)
    b2 = 0.3
    mid = F.conv2d(x.float(), wt.float(), b, padding=1).relu()
    want = (torch.einsum("nchw,c->nhw", mid, w2) + b2).relu()
    xn = x.permute(0, 2, 3, 1).contiguous()
    wk = wt.permute(0, 2, 3, 1).reshape(32, 9 * C).contiguous()
    out = torch.full((n, H, W), float("nan"), device="cuda")
    L.conv3x3(xn, wk, L.make_epilogue(bias=b, out_f32=out, w2=w2, b2=b2, store_mode=L.ST_HEADDOT), n, H, W, C, 32)
    _close(out, want, 1e-2)


def test_gemm_throughput_report(L, capsys):
    """Not an assertion on speed: prints the achieved TFLOP/s of the encoder-sized GEMMs."""
    M = 32 * 1370
    for N, K in [(3072, 1024), (1024, 1024), (4096, 1024), (1024, 4096)]:
        a = _rand(M, K, seed=1).bfloat16()
        w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        epi = L.make_epilogue(out=out, ldo=N)
        for _ in range(3):
            L.gemm(a, w, epi)
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(10):
            L.gemm(a, w, epi)
        t1.record()
        torch.cuda.synchronize()
        ms = t0.elapsed_time(t1) / 10
        ref0, ref1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3):
            torch.matmul(a, w.T)
        ref0.record()
        for _ in range(10):
            torch.matmul(a, w.T)
        ref1.record()
        torch.cuda.synchronize()
        ms_ref = ref0.elapsed_time(ref1) / 10
        with capsys.disabled():
            print(f"\n[gemm] M={M} N={N} K={K}: {ms:.3f} ms = {2 * M * N * K / ms / 1e9:.0f} TFLOP/s "
                  f"(cuBLAS {ms_ref:.3f} ms = {2 * M * N * K / ms_ref / 1e9:.0f} TFLOP/s)")
        _close(out, a.float() @ w.float().T)


def _guarded(rows, cols, fill=float("nan")):
    """bf16 [rows, cols] view inside a buffer with sentinel bands on both sides (scattered stores must stay inside)."""
    pad = 4096
    buf = torch.full((pad + rows * cols + pad,), 7.0, device="cuda", dtype=torch.bfloat16)
    view = buf[pad:pad + rows * cols].view(rows, cols)
    view.fill_(fill)
    return buf, view, pad


def _guards_intact(buf, pad):
    return bool((buf[:pad] == 7.0).all()) and bool((buf[-pad:] == 7.0).all())


@pytest.mark.parametrize("M,N,K,pair", [(999, 256, 256, 0), (1370, 1024, 1024, 0), (130, 96, 192, 0), (5000, 64, 64, 0),
                                        (700, 512, 128, 2), (43808, 256, 256, 0), (257, 32, 64, 1)])
@pytest.mark.parametrize("combo", ["r1", "r2", "r1r2", "relu_only", "r1_relu", "r1r2_relu", "relu_no_out"])
def test_gemm_cooperative_box_epilogue(L, M, N, K, pair, combo):
    """bf16 residuals and / or a ReLU'd second output through the cooperative-box epilogue (MODE 5): every combination,
    ragged M, N smaller than a tile, 1-CTA and CTA-pair kernels; outputs NaN-filled between sentinel bands."""
    a = _rand(M, K, seed=1).bfloat16()
    w = _rand(N, K, scale=K ** -0.5, seed=2).bfloat16()
    b = _rand(N, seed=3)
    r1 = _rand(M, N, seed=4).bfloat16() if "r1" in combo else None
    r2 = _rand(M, N, seed=5).bfloat16() if "r2" in combo else None
    want = a.float() @ w.float().T + b
    if r1 is not None:
        want = want + r1.float()
    if r2 is not None:
        want = want + r2.float()
    bo, out, pad = _guarded(M, N)
    br, out_relu, _ = _guarded(M, N)
    use_out = combo != "relu_no_out"
    use_relu = "relu" in combo
    L.gemm(a, w, L.make_epilogue(bias=b, res1=r1, res2=r2, out=out if use_out else None,
                                 out_relu=out_relu if use_relu else None, ldo=N, cta_pair=pair))
    torch.cuda.synchronize()
    if use_out:
        _close(out, want)
    if use_relu:
        _close(out_relu, want.relu())
    assert _guards_intact(bo, pad) and _guards_intact(br, pad)


@pytest.mark.parametrize("n,H,W,C,Cout", [(2, 19, 19, 64, 64), (1, 37, 37, 128, 256), (2, 74, 74, 256, 256), (1, 5, 200, 192, 128),
                                          (1, 148, 148, 64, 32)])
@pytest.mark.parametrize("combo", ["r1r2_relu", "r2", "relu_only"])
def test_conv3x3_cooperative_box_epilogue(L, n, H, W, C, Cout, combo):
    """The FeatureFusion / ResidualConvUnit epilogues (util/blocks.py:78-91, 146-160) on maps whose sizes are not
    multiples of the pixel tile, incl. the CTA-pair halo mode (Cout 128 / 256)."""
    x = _rand(n, C, H, W, seed=1).bfloat16()
    wt = _rand(Cout, C, 3, 3, scale=(9 * C) ** -0.5, seed=2).bfloat16()
    b = _rand(Cout, seed=3)
    r1 = _rand(n, H, W, Cout, seed=4).bfloat16() if "r1" in combo else None
    r2 = _rand(n, H, W, Cout, seed=5).bfloat16() if "r2" in combo else None
    want = F.conv2d(x.float(), wt.float(), b, padding=1).permute(0, 2, 3, 1)
    if r1 is not None:
        want = want + r1.float()
    if r2 is not None:
        want = want + r2.float()
    xn = x.permute(0, 2, 3, 1).contiguous()
    wk = wt.permute(0, 2, 3, 1).reshape(Cout, 9 * C).contiguous()
    bo, out, pad = _guarded(n * H * W, Cout)
    br, out_relu, _ = _guarded(n * H * W, Cout)
    use_relu = "relu" in combo
    L.conv3x3(xn, wk, L.make_epilogue(bias=b, res1=r1, res2=r2, out=out, out_relu=out_relu if use_relu else None, ldo=Cout),
              n, H, W, C, Cout)
    torch.cuda.synchronize()
    _close(out.view(n, H, W, Cout), want)
    if use_relu:
        _close(out_relu.view(n, H, W, Cout), want.relu())
    assert _guards_intact(bo, pad) and _guards_intact(br, pad)

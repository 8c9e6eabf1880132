"""vda_groupnorm_nhwc_static: the allocation-free GroupNorm entry point (scratch = a file-scope `static __device__`
array inside the library, include/vda_b200.h).  No counterpart in the reference.
Collected last on purpose (file name).

The GPU test calls the new entry point through the C ABI twice in a row (the array keeps the first call's contents:
the second call checks that what must start at zero is cleared) and compares with vda_groupnorm_nhwc on caller-owned
scratch, and with torch's group_norm at the tolerance tests/test_kernels_gpu.py uses (1e-2 of max |want|)."""
import pytest
import torch
import torch.nn.functional as F


def _rand(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).cuda()


@pytest.fixture(scope="module")
def L():
    import vda_b200.lib as lib

    lib.init()
    return lib


@pytest.mark.gpu
@pytest.mark.parametrize("n,h,w,C", [(32, 37, 37, 256), (3, 19, 19, 1024), (64, 5, 7, 192)])
def test_groupnorm_static_matches_caller_scratch_twice(L, n, h, w, C):
    g, b = _rand(C, seed=2), _rand(C, seed=3)
    outs = []
    for call in range(2):      # different inputs per call: stale sums of call 0 would show up in call 1
        x = (_rand(n, C, h, w, scale=2.0 + call, seed=10 + call) + 1.0 - 3.0 * call).bfloat16()
        xn = x.permute(0, 2, 3, 1).contiguous()
        want = F.group_norm(x.float(), 32, g, b, 1e-6).permute(0, 2, 3, 1)
        ref = torch.full((n, h, w, C), float("nan"), device="cuda", dtype=torch.bfloat16)
        got = torch.full_like(ref, float("nan"))
        ws = torch.empty(96 * n, device="cuda", dtype=torch.float64)
        L.groupnorm_nhwc(xn, g, b, ref, ws, n, h * w, C, 1e-6)
        torch.cuda.synchronize()
        alloc0, reserved0 = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
        free0 = torch.cuda.mem_get_info()[0] if call else None      # call 0 may load the module lazily
        L.groupnorm_nhwc_static(xn, g, b, got, n, h * w, C, 1e-6)
        torch.cuda.synchronize()
        assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == (alloc0, reserved0)
        if call:
            assert torch.cuda.mem_get_info()[0] >= free0, "device memory was allocated inside the call"
        tol = 1e-2 * want.abs().max().item()
        assert torch.isfinite(got.float()).all()
        assert (got.float() - ref.float()).abs().max().item() <= tol      # fp64 atomics: order of the sums is not fixed
        assert (got.float() - want).abs().max().item() <= tol
        outs.append(got)
    assert not torch.equal(outs[0], outs[1])


@pytest.mark.gpu
def test_groupnorm_static_split_mode(L):
    """FP32-accuracy mode layout ([hi | lo | hi] in, split out) through the static scratch."""
    n, h, w, C = 2, 10, 10, 64
    x = _rand(n, h, w, C, scale=2.0, seed=5) + 0.5
    hi = x.bfloat16()
    lo = (x - hi.float()).bfloat16()
    xs = torch.cat([hi, lo, hi], dim=-1).contiguous()
    g, b = _rand(C, seed=2), _rand(C, seed=3)
    ref = torch.empty(n, h, w, 3 * C, device="cuda", dtype=torch.bfloat16)
    got = torch.empty_like(ref)
    ws = torch.empty(96 * n, device="cuda", dtype=torch.float64)
    L.groupnorm_nhwc(xs, g, b, ref, ws, n, h * w, C, 1e-6, ldx=3 * C, split=C)
    for _ in range(2):
        got.fill_(float("nan"))
        L.groupnorm_nhwc_static(xs, g, b, got, n, h * w, C, 1e-6, ldx=3 * C, split=C)
        torch.cuda.synchronize()
        v = got[..., :C].float() + got[..., C:2 * C].float()
        r = ref[..., :C].float() + ref[..., C:2 * C].float()
        assert (v - r).abs().max().item() <= 1e-5 * r.abs().max().item()


def test_groupnorm_static_rejects_sizes_beyond_the_scratch():
    """Argument check happens before any CUDA call, so it runs without a GPU too."""
    import vda_b200.lib as lib

    h = lib.load()
    rc = h.vda_groupnorm_nhwc_static(None, 64, None, None, None, lib.GROUPNORM_STATIC_MAX_N + 1, 16, 64, 1e-6, 0, None)
    assert rc == -8 and b"exceed the static scratch" in h.vda_last_error()
    with pytest.raises(lib.VdaError):
        lib.check(rc)

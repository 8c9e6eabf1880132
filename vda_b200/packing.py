"""One-time repacking of reference-layout fp32 parameters into the kernel layouts.

Everything here runs once after load_state_dict()/.to(device) (SURVEY.md 8b "Lifecycle"), never per
window.  Layout rules:
  * every GEMM weight is bf16 [N, K] with K contiguous; N is padded to a multiple of 16 and, when the
    output feeds a 3x3 implicit GEMM as input channels, to a multiple of 64 (zero rows/bias);
  * 3x3 conv weights are [Cout, (kh, kw, cin_padded)];
  * ConvTranspose(k=s) weights are [(kh, kw, cout_padded), cin];
  * GEGLU projection rows are interleaved in 16-row value/gate groups (gemm.cu epilogue).
"""
import contextlib

import torch

# FP32-accuracy mode (engine.Engine(fp32=True)): every weight is packed [Wh | Wh | Wl] along K (per
# 3x3 tap for the implicit-GEMM convs), Wh = bf16(W), Wl = bf16(W - Wh), to be multiplied with split
# activations [hi | lo | hi]: x.W ~= hi.Wh + lo.Wh + hi.Wl (include/vda_b200.h, vda_epilogue::split).
SPLIT = False


@contextlib.contextmanager
def split_mode(on):
    global SPLIT
    old, SPLIT = SPLIT, bool(on)
    try:
        yield
    finally:
        SPLIT = old


def to_kernel(w):
    """f32 [..., K] -> bf16 [..., K], or [..., 3K] = [Wh | Wh | Wl] in split mode."""
    hi = w.bfloat16()
    if not SPLIT:
        return hi.contiguous()
    lo = (w - hi.float()).bfloat16()
    return torch.cat([hi, hi, lo], dim=-1).contiguous()


def _pad_to(n, m):
    return (n + m - 1) // m * m


def linear(w, b=None, n_pad=16, device=None):
    """nn.Linear / 1x1 conv weight [N, K(,1,1)] -> (bf16 [Np, K], f32 bias [Np] or None)."""
    w = w.detach().reshape(w.shape[0], -1).float()
    n, k = w.shape
    npad = _pad_to(n, n_pad)
    wk = torch.zeros(npad, k, dtype=torch.float32, device=device or w.device)
    wk[:n] = w.to(wk.device)
    bk = None
    if b is not None:
        bk = torch.zeros(npad, dtype=torch.float32, device=wk.device)
        bk[:n] = b.detach().float().to(wk.device)
    return to_kernel(wk), bk


def conv3x3(w, b=None, cin_pad=64, cout_pad=16, device=None):
    """Conv2d 3x3 weight [Co, Ci, 3, 3] -> (bf16 [Cop, 9*Cip], f32 bias [Cop] or None), k=(kh,kw,ci)."""
    w = w.detach().float()
    co, ci = w.shape[:2]
    cop, cip = _pad_to(co, cout_pad), _pad_to(ci, cin_pad)
    dev = device or w.device
    wk = torch.zeros(cop, 3, 3, cip, dtype=torch.float32, device=dev)
    wk[:co, :, :, :ci] = w.permute(0, 2, 3, 1).to(dev)
    bk = None
    if b is not None:
        bk = torch.zeros(cop, dtype=torch.float32, device=dev)
        bk[:co] = b.detach().float().to(dev)
    return to_kernel(wk.reshape(cop, 9, cip)).reshape(cop, -1).contiguous(), bk


def conv_transpose(w, b, cout_pad=64, device=None):
    """ConvTranspose2d(k=s) weight [Ci, Co, s, s] -> (bf16 [s*s*Cop, Ci], f32 bias [s*s*Cop], s, Cop)."""
    w = w.detach().float()
    ci, co, s, _ = w.shape
    cop = _pad_to(co, cout_pad)
    dev = device or w.device
    wk = torch.zeros(s, s, cop, ci, dtype=torch.float32, device=dev)
    wk[:, :, :co] = w.permute(2, 3, 1, 0).to(dev)
    bk = torch.zeros(s, s, cop, dtype=torch.float32, device=dev)
    bk[:, :, :co] = b.detach().float().to(dev)
    return to_kernel(wk.reshape(s * s * cop, ci)), bk.reshape(-1).contiguous(), s, cop


def geglu(w, b, device=None):
    """GEGLU proj [2*inner, C] (value rows then gate rows) -> rows interleaved in groups of 16."""
    n = w.shape[0]
    inner = n // 2
    assert inner % 16 == 0
    idx = torch.arange(n).view(2, inner // 16, 16).permute(1, 0, 2).reshape(-1)
    dev = device or w.device
    return (to_kernel(w.detach().float()[idx].to(dev)),
            b.detach().float()[idx].to(dev).contiguous())


def vec(t, device=None, n_pad=1):
    t = t.detach().float().reshape(-1)
    out = torch.zeros(_pad_to(t.numel(), n_pad), dtype=torch.float32, device=device or t.device)
    out[: t.numel()] = t.to(out.device)
    return out

"""Phase clocks of the persistent ping-pong attention kernel (PROF build of attn_fwd_v3_kernel), per KV tile and
stream, vitl block (32 frames x 16 heads x S=1370).  VDA_ATTN_PP=0: free-running streams."""
# needs the experiment build:  export VDA_NVCC_EXTRA=-DVDA_ATTN_EXPERIMENTS VDA_LIB_NAME=libvda_b200_exp.so
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import vda_b200.lib as L

L.init()
T, S, heads = 32, 1370, 16
D = heads * 64
spad = (S + 7) // 8 * 8
g = torch.Generator(device="cuda").manual_seed(0)
qk = torch.randn(T * S, 2 * D, device="cuda", generator=g).bfloat16()
vt = torch.randn(T * heads, 64, spad, device="cuda", generator=g).bfloat16()
ao = torch.empty(T * S, D, device="cuda", dtype=torch.bfloat16)
qt = (S + 127) // 128
nslots = 2 * L.load().vda_num_sms()
prof = torch.zeros(nslots * 4, 8, device="cuda", dtype=torch.int64)
lib = L.load()
lib.vda_debug_attn_prof.argtypes = [C.c_void_p]
lib.vda_debug_attn_prof.restype = None
for _ in range(3):
    L.flash_attn(qk, vt, ao, T, S, heads, spad)
lib.vda_debug_attn_prof(L.ptr(prof))
L.flash_attn(qk, vt, ao, T, S, heads, spad)
torch.cuda.synchronize()
lib.vda_debug_attn_prof(None)
p = prof.double()
tiles = p[:, 7:8] * qt
names = ["wait S_g", "ld S + row max", "wait for the other stream", "exp2 + pack", "wait PV_(g-1) + st P + p_full"]
for i, n in enumerate(names):
    v = p[:, i:i + 1] / tiles
    print(f"{n:32s} {v.mean().item():8.1f} clk per KV tile per stream (p10 {v.quantile(0.1).item():7.1f}, p90 {v.quantile(0.9).item():7.1f})")
print(f"{'epilogue per item':32s} {(p[:, 5] / p[:, 7]).mean().item():8.1f} clk")
print(f"{'whole stream / tiles':32s} {(p[:, 6:7] / tiles).mean().item():8.1f} clk per KV tile; stream total max {p[:, 6].max().item():.0f} min {p[:, 6].min().item():.0f} clk; items per stream {p[:, 7].min().item():.0f}..{p[:, 7].max().item():.0f}")

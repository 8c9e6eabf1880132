"""Where a softmax warp of the spatial attention kernel spends its clocks (PROF build of attn_fwd_v2_kernel):
per KV tile, averaged over all CTAs of a vitl block (32 frames x 16 heads x S=1370)."""
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
prof = torch.zeros(T * heads * qt * 4, 8, device="cuda", dtype=torch.int64)
lib = L.load()
lib.vda_debug_attn_prof.argtypes = [C.c_void_p]
lib.vda_debug_attn_prof.restype = None
for _ in range(3):
    L.flash_attn(qk, vt, ao, T, S, heads, spad)
lib.vda_debug_attn_prof(L.ptr(prof))
L.flash_attn(qk, vt, ao, T, S, heads, spad)
torch.cuda.synchronize()
lib.vda_debug_attn_prof(None)
p = prof.double() / qt           # per KV tile
names = ["wait S_j", "tcgen05.ld S + s_free", "row max", "exp2 + pack", "wait PV_(j-1)", "tcgen05.st P + p_full", "loop total"]
m = p.mean(dim=0)
for i, n in enumerate(names):
    print(f"{n:28s} {m[i].item():8.1f} clk per KV tile per CTA   (p10 {p[:, i].quantile(0.1).item():7.1f}, p90 {p[:, i].quantile(0.9).item():7.1f})")
print(f"sum of phases {m[:6].sum().item():.1f}")

"""Streaming temporal attention kernel (vda_temporal_attn_stream) per motion-module shape of vits / vitb / vitl at
518x518: us per launch and effective GB/s of cache traffic, 2 vs 4 tokens per thread."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import vda_b200.lib as L

L.init()
shapes = {"vits": [(1369, 192), (361, 384), (1369, 64), (5476, 64)], "vitb": [(1369, 384), (361, 768), (1369, 128), (5476, 128)],
          "vitl": [(1369, 1024), (361, 1024), (1369, 256), (5476, 256)]}
cap, Tk = 44, 32
for arch, lst in shapes.items():
    for hw, C in lst:
        qkv = torch.randn(hw, 3 * C, device="cuda").bfloat16()
        kc = torch.randn(hw, cap, C, device="cuda").bfloat16()
        vc = torch.randn(hw, cap, C, device="cuda").bfloat16()
        pe = [torch.randn(32, C, device="cuda") for _ in range(3)]
        slots = torch.arange(32, dtype=torch.int32, device="cuda")
        out = torch.empty(hw, C, device="cuda", dtype=torch.bfloat16)
        flush = torch.empty(64 << 20, device="cuda", dtype=torch.float32)
        res = []
        for mode in ("tb2", "tb4"):
            os.environ["VDA_TSTREAM_TB"] = mode[2:]
            fn = lambda: L.temporal_attn_stream(qkv, qkv[:, C:], qkv[:, 2 * C:], 3 * C, kc, vc, slots, *pe, out, C, hw, 8, C // 8, Tk)
            fn()
            best = 1e9
            for _ in range(5):
                flush.sum()        # 256 MB read > L2: the caches come from HBM as in a real frame (clean lines)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record(); torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1))
            res.append(best)
        mb = 2 * Tk * hw * C * 2 / 1e6
        print(f"{arch} hw={hw:5d} C={C:5d}: TB=2 {res[0] * 1e3:6.1f} us, TB=4 {res[1] * 1e3:6.1f} us (incl. ~35 us of launch overhead)  ({mb:6.1f} MB of cache)", flush=True)

"""Where a streaming frame's time goes (config 5): after the cache is in steady state, one frame is run eagerly with
CUDA events around every instrumented launch (vda_b200.lib.PROFILE) and the launches are grouped by kind and shape.
    python scripts/prof_stream.py [--arch vitl]"""
import argparse
import collections
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import vda_b200.lib as L
from vda_b200.video_depth import MODEL_CONFIGS, randomize_zero_init_
from video_depth_anything.video_depth_stream import VideoDepthAnything

ap = argparse.ArgumentParser()
ap.add_argument("--arch", default="vitl")
ap.add_argument("--size", type=int, default=518)
a = ap.parse_args()
torch.manual_seed(0)
model = randomize_zero_init_(VideoDepthAnything(**MODEL_CONFIGS[a.arch])).cuda().eval()
rng = np.random.default_rng(0)
frames = rng.integers(0, 256, (80, a.size, a.size, 3), dtype=np.uint8)
lat = []
for i in range(60):
    t0 = time.perf_counter()
    model.infer_video_depth_one(frames[i], input_size=518, device="cuda")
    lat.append((time.perf_counter() - t0) * 1e3)
print(f"{a.arch}: graph-replay latency p50 {np.percentile(lat[45:], 50):.3f} ms")
eng = model.engine()
eng.use_graph = False
for i in range(60, 64):
    t0 = time.perf_counter()
    model.infer_video_depth_one(frames[i], input_size=518, device="cuda")
    print(f"eager frame: {(time.perf_counter() - t0) * 1e3:.3f} ms wall")
L.PROFILE, L.NOTES = [], []
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
model.infer_video_depth_one(frames[64], input_size=518, device="cuda")
e1.record()
torch.cuda.synchronize()
prof, notes = L.PROFILE, L.NOTES
L.PROFILE = L.NOTES = None
print(f"instrumented eager frame: {e0.elapsed_time(e1):.3f} ms on the device, {len(prof)} instrumented launches")
by = collections.OrderedDict()
for (kind, flops, s, e), note in zip(prof, notes):
    k = by.setdefault(note or kind, [0, 0.0, 0.0])
    k[0] += 1
    k[1] += s.elapsed_time(e)
    k[2] += flops
tot = sum(v[1] for v in by.values())
print(f"sum of instrumented launches {tot:.3f} ms")
for k, v in sorted(by.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{v[1]:8.3f} ms  n={v[0]:3d}  {v[1] / v[0] * 1e3:7.1f} us each  {v[2] / max(v[1], 1e-9) / 1e9:7.0f} TF/s|GB/s  {k}")

"""BASELINE configs[4]: streaming mode latency per frame (video_depth_stream.py), vits 518x518 by default.
Per-frame wall time of infer_video_depth_one: uint8 frame in, float32 depth out on the host."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from vda_b200.video_depth import MODEL_CONFIGS, randomize_zero_init_
from video_depth_anything.video_depth_stream import VideoDepthAnything

ap = argparse.ArgumentParser()
ap.add_argument("--arch", default="vits")
ap.add_argument("--frames", type=int, default=150)
ap.add_argument("--size", type=int, default=518)
ap.add_argument("--no-graph", action="store_true")
a = ap.parse_args()
torch.manual_seed(0)
cfg = {k: v for k, v in MODEL_CONFIGS[a.arch].items()}
model = randomize_zero_init_(VideoDepthAnything(**cfg)).cuda().eval()
if a.no_graph:
    model.engine().use_graph = False
rng = np.random.default_rng(0)
frames = rng.integers(0, 256, (a.frames, a.size, a.size, 3), dtype=np.uint8)
lat = []
for i in range(a.frames):
    t0 = time.perf_counter()
    d = model.infer_video_depth_one(frames[i], input_size=518, device="cuda")
    lat.append((time.perf_counter() - t0) * 1e3)
steady = np.array(lat[45:])   # past the cache steady state (42-entry list full)
print(json.dumps({"workload": f"streaming {a.arch} {a.size}x{a.size}, {a.frames} frames, graph={not a.no_graph}",
                  "p50_ms": float(np.percentile(steady, 50)), "p99_ms": float(np.percentile(steady, 99)),
                  "mean_ms": float(steady.mean()), "first_frames_ms": lat[:4], "fps": 1e3 / float(steady.mean()),
                  "depth_range": [float(d.min()), float(d.max())]}))

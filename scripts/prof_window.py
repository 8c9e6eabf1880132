"""One warm-up + N window forwards of the CUDA path, for ncu (no timing printed from here is a bench value)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from vda_b200.video_depth import MODEL_CONFIGS, randomize_zero_init_
from video_depth_anything.video_depth import VideoDepthAnything

ap = argparse.ArgumentParser()
ap.add_argument("--arch", default="vitl")
ap.add_argument("--frames", type=int, default=32)
ap.add_argument("--size", type=int, default=518)
ap.add_argument("--iters", type=int, default=1)
a = ap.parse_args()
torch.manual_seed(0)
torch.set_grad_enabled(False)
m = randomize_zero_init_(VideoDepthAnything(**MODEL_CONFIGS[a.arch])).cuda().eval()
x = torch.randn(a.frames, 3, a.size, a.size, device="cuda")
eng = m.engine()
for _ in range(1 + a.iters):
    d = eng.forward(x)
torch.cuda.synchronize()
print("ok", float(d.mean()))

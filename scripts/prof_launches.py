"""Per-launch CUDA-event times of the tensor-core launches of one window (eager), with shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vda_b200.lib as L
from vda_b200.video_depth import MODEL_CONFIGS, randomize_zero_init_
from video_depth_anything.video_depth import VideoDepthAnything

arch = sys.argv[1] if len(sys.argv) > 1 else "vitl"
fp32 = len(sys.argv) > 2 and sys.argv[2] == "fp32"   # FP32-accuracy mode (split-bf16 products)
torch.manual_seed(0); torch.set_grad_enabled(False)
m = randomize_zero_init_(VideoDepthAnything(**MODEL_CONFIGS[arch])).cuda().eval()
x = torch.randn(32, 3, 518, 518, device="cuda")
eng = m.engine(fp32)
for _ in range(2): eng.forward(x)
best = None
for rep in range(3):
    L.PROFILE, L.NOTES = [], []
    eng.forward(x); torch.cuda.synchronize()
    t = [s.elapsed_time(e) for _, _, s, e in L.PROFILE]
    best = t if best is None else [min(a, b) for a, b in zip(best, t)]
    prof, notes = L.PROFILE, L.NOTES
L.PROFILE = L.NOTES = None
agg = {}
for (kind, fl, _, _), note, ms in zip(prof, notes, best):
    k = agg.setdefault(note or kind, [0, 0.0, 0.0]); k[0] += 1; k[1] += ms; k[2] += fl
print(f"{'ms':>8} {'n':>3} {'TF/s':>6}  launch")
for note, (n, ms, fl) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{ms:8.3f} {n:3d} {fl / ms / 1e9:6.0f}  {note}")
print(f"total {sum(best):.2f} ms over {len(best)} tensor-core launches")

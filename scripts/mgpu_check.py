"""torchrun --nproc-per-node N scripts/mgpu_check.py : sharded infer_video_depth == single-GPU result."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from vda_b200.video_depth import (MODEL_CONFIGS, align_windows_device, randomize_zero_init_, window_frame_indices)
from video_depth_anything.video_depth import VideoDepthAnything

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
torch.manual_seed(0)
m = randomize_zero_init_(VideoDepthAnything(**MODEL_CONFIGS["vits"])).to(f"cuda:{local}").eval()
n, h, w = 120, 90, 120
rng = np.random.default_rng(0)
yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
frames = np.stack([np.clip(127 + 100 * np.sin((xx + 2 * t) / 9)[..., None] * np.cos((yy - t) / 7)[..., None]
                           + rng.uniform(-20, 20, (h, w, 3)), 0, 255).astype(np.uint8) for t in range(n)])
got, fps = m.infer_video_depth(frames, 30, input_size=112, device="cuda")
n_win = len(window_frame_indices(n))
raw = m.infer_video_depth_device(frames, 112, "cuda", windows=range(n_win), return_raw=True)
want = align_windows_device(raw, False)[:n].cpu().numpy()
err = float(np.abs(got - want).max() / np.abs(want).max())
print(f"rank {dist.get_rank()}/{dist.get_world_size()}: windows {n_win}, max rel diff vs single-GPU {err:.2e}, "
      f"range {want.min():.3f}..{want.max():.3f}", flush=True)
assert got.shape == (n, h, w) and err < 1e-4
dist.barrier()
dist.destroy_process_group()

"""BASELINE configs[2]/[3]: whole-video throughput of infer_video_depth (pre-processing, window forwards,
resize to source, alignment scan, D2H), one process per GPU under torchrun.  Prints video frames/s.

    python scripts/video_bench.py [--arch vitl] [--frames 1024] [--height 518] [--width 518] [--metric]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from vda_b200.video_depth import MODEL_CONFIGS, randomize_zero_init_, window_frame_indices
from video_depth_anything.video_depth import VideoDepthAnything

ap = argparse.ArgumentParser()
ap.add_argument("--arch", default="vitl")
ap.add_argument("--frames", type=int, default=1024)
ap.add_argument("--height", type=int, default=518)
ap.add_argument("--width", type=int, default=518)
ap.add_argument("--input-size", type=int, default=518)
ap.add_argument("--metric", action="store_true")
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--result-rank", default="all", help="all: every rank returns the full result; 0: only rank 0 does; sharded: every rank keeps its own frames; both: all, then 0")
a = ap.parse_args()

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
torch.manual_seed(0)
cfg = dict(MODEL_CONFIGS[a.arch], metric=a.metric)
model = randomize_zero_init_(VideoDepthAnything(**cfg)).to(f"cuda:{local}").eval()

# smooth pattern translating over time + noise, built in chunks to bound host memory
rng = np.random.default_rng(0)
yy, xx = np.mgrid[0:a.height, 0:a.width].astype(np.float32)
frames = np.empty((a.frames, a.height, a.width, 3), np.uint8)
for t in range(a.frames):
    base = 127 + 90 * np.sin((xx + 3 * t) / 37) * np.cos((yy - 2 * t) / 29)
    frames[t] = np.clip(base[..., None] + rng.integers(-12, 13, (a.height, a.width, 3)), 0, 255).astype(np.uint8)

from vda_b200 import parallel

modes = ["all", "0"] if a.result_rank == "both" else [a.result_rank]
for mode in modes:
    if mode == "sharded":
        parallel.set_result_sharded()
    else:
        parallel.set_result_rank(None if mode == "all" else int(mode))
    times = []
    for rep in range(a.reps + 1):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        depth, _ = model.infer_video_depth(frames, 24, input_size=a.input_size, device="cuda")
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        times.append(time.perf_counter() - t0)
        tr = model.last_timing.get("trace")
        if tr is not None and rep == a.reps and rank in (0, world // 2, world - 1):   # VDA_VIDEO_TRACE=1: per-round times
            for j, r in enumerate(tr):
                print(f"trace mode={mode} rank {rank} round {j}: host " + " ".join(f"{k}={v:.1f}" for k, v in r["host_ms"].items())
                      + " | device " + " ".join(f"{k}={v:.1f}" for k, v in r["device_ms"].items()), flush=True)
            print(f"trace mode={mode} rank {rank} total {times[-1] * 1e3:.1f} ms, device {model.last_timing['seconds_device'] * 1e3:.1f} ms", flush=True)
    best = min(times[1:])
    if rank == 0:
        n_win = len(window_frame_indices(a.frames))
        print(json.dumps({"workload": f"infer_video_depth {a.arch} {a.frames} frames {a.height}x{a.width} metric={a.metric}",
                          "n_gpus": world, "result_rank": mode, "windows": n_win, "seconds": best,
                          "video_frames_per_s": a.frames / best, "window_frames_per_s": 32 * n_win / best, "all_runs_s": times,
                          "depth_shape": list(depth.shape), "depth_range": [float(depth.min()), float(depth.max())]}), flush=True)
if world > 1:
    dist.destroy_process_group()

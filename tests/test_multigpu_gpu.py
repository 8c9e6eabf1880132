"""Sharded infer_video_depth over NCCL equals the single-GPU result (needs >= 2 GPUs; skipped otherwise)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on the box")
def test_sharded_driver_matches_single_gpu():
    import socket

    n = min(torch.cuda.device_count(), 4)
    with socket.socket() as sk:            # a free rendezvous port: a fixed one can still be held by an earlier torchrun
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "scripts", "mgpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    print(r.stdout[-2000:])
    if r.returncode != 0:        # the whole stderr, not a tail: exit-time NCCL warnings bury the Python traceback
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "mgpu_check_failure.log"), "w") as f:
            f.write(r.stdout + "\n==== stderr ====\n" + r.stderr)
    tb = r.stderr[r.stderr.find("Traceback"):] if "Traceback" in r.stderr else r.stderr
    assert r.returncode == 0, (r.stdout[-1500:] + "\n" + tb[:4000])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on the box")
def test_two_devices_in_one_process():
    """Per-device launch state (kernel attributes, SM count): a model on cuda:1 after one on cuda:0 in the same
    process -- the single-process multi-device mode SURVEY.md 8(e) names as the run.py alternative."""
    import numpy as np

    from oracle import make_golden as G
    from oracle import refload
    from oracle import vda_oracle as O
    from video_depth_anything.video_depth import VideoDepthAnything

    arch, T, H, W = "vits", 4, 70, 98
    cfg = O.ARCH[arch]
    sd = refload.seeded_state_dict(arch, 0)
    x = G.window_input(T, H, W)
    gold = np.load(os.path.join(G.GOLDEN_DIR, "fwd_vits_t4_70x98.npz"))["depth"]
    outs = []
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        m = VideoDepthAnything(encoder=arch, features=cfg["features"], out_channels=cfg["out_channels"])
        m.load_state_dict(sd, strict=True)
        m = m.to(dev).eval()
        with torch.no_grad(), torch.cuda.device(dev):
            outs.append(m(x.to(dev)).cpu().numpy())
    for o in outs:
        msk = gold > 1e-6
        assert float(np.mean(np.abs(o[msk] - gold[msk]) / gold[msk])) <= 1e-2
    # same kernels, same input, another device: equal up to the order of the GroupNorm / split-K atomics (a last-bit
    # difference that a bf16 rounding can amplify at single pixels), so mean-relative tight, max loose
    d = np.abs(outs[0] - outs[1])
    assert d.mean() <= 1e-4 * np.abs(outs[0]).mean() and d.max() <= 2e-2 * np.abs(outs[0]).max()

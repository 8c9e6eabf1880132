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
    assert r.returncode == 0, (r.stdout[-1500:] + "\n" + r.stderr[-3000:])

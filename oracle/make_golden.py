"""ORACLE support (test infrastructure): generates tests/golden/*.npz by running the UNMODIFIED
reference (/root/reference, build container only) on seeded weights and inputs.

    python -m oracle.make_golden

The fixtures pin oracle/vda_oracle.py (tests/test_oracle.py) and, on the GPU box where the reference
is absent, stand in for it.  Inputs are regenerated from seeds by golden_inputs(), so only outputs
are stored.
"""
import os

import numpy as np
import torch

from . import refload
from .vda_oracle import ARCH

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

FORWARD_CASES = {  # name -> (arch, T, H, W)
    "fwd_vits_t4_70x98": ("vits", 4, 70, 98),
    "fwd_vits_t32_56x56": ("vits", 32, 56, 56),
    "fwd_vitb_t2_42x56": ("vitb", 2, 42, 56),
    "fwd_vitl_t2_28x42": ("vitl", 2, 28, 42),
}
VIDEO_CASE = dict(arch="vits", n=46, h=44, w=60, input_size=56)


def window_input(T, H, W, seed=1):
    return torch.randn(1, T, 3, H, W, generator=torch.Generator().manual_seed(seed))


def video_frames(n, h, w, seed=0):
    """Smooth pattern drifting over time + noise, uint8 RGB [n,h,w,3]."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    out = np.empty((n, h, w, 3), np.uint8)
    for t in range(n):
        base = [127 + 100 * np.sin((xx + 2 * t) / (7 + 3 * c)) * np.cos((yy - t) / (5 + 2 * c)) for c in range(3)]
        img = np.stack(base, -1) + rng.uniform(-20, 20, (h, w, 3))
        out[t] = np.clip(img, 0, 255).astype(np.uint8)
    return out


UTIL_SHAPE = (2, 37, 53)
NORMALIZE_CASES = [(0.0, 1.0, False), (0.1, 10.0, True), (5.0, 0.5, False), (2.0, 2.0, True)]  # near, far, invert


def util_inputs(seed=0):
    """pred, target, mask (float32 [2,37,53]; mask has fractional weights and zeros) and 8+8 frames to cross-fade."""
    rng = np.random.default_rng(seed)
    pred = rng.uniform(0.2, 7.5, UTIL_SHAPE).astype(np.float32)
    target = (1.7 * pred - 0.3 + rng.normal(0, 0.05, UTIL_SHAPE)).astype(np.float32)
    mask = (rng.uniform(0, 1, UTIL_SHAPE) > 0.3).astype(np.float32) * rng.choice([1.0, 0.5], UTIL_SHAPE).astype(np.float32)
    pre = [rng.normal(0, 2, UTIL_SHAPE[1:]).astype(np.float32) for _ in range(8)]
    post = [rng.normal(0, 2, UTIL_SHAPE[1:]).astype(np.float32) for _ in range(8)]
    depth = rng.uniform(0.0, 80.0, (3,) + UTIL_SHAPE[1:]).astype(np.float32)
    depth[2] = 4.25   # constant frame (engine.py:1123-1124)
    return pred, target, mask, pre, post, depth


def make_util_golden():
    """utils/util.py:16-74 and external_engine/engine.py:1099-1141 outputs of the unmodified reference."""
    csas, gif, nd = refload.import_reference_utils()
    pred, target, mask, pre, post, depth = util_inputs()
    out = {}
    out["full"] = np.array(csas(pred, target, mask), np.float64)
    out["full_ones"] = np.array(csas(pred, target, np.ones_like(mask)), np.float64)
    out["scale_only"] = np.array(csas(pred, target, mask, scale_only=True), np.float64)
    out["blend"] = np.stack(gif(pre, post)).astype(np.float32)
    for i, (near, far, inv) in enumerate(NORMALIZE_CASES):
        out[f"norm{i}"] = np.stack([nd(depth[f], near, far, inv) for f in range(depth.shape[0])]).astype(np.float32)
    np.savez_compressed(os.path.join(GOLDEN_DIR, "util_cases.npz"), **out)
    print("util", {k: v.shape for k, v in out.items()})


def main():
    assert refload.reference_available(), "needs /root/reference (build container)"
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    make_util_golden()
    if os.environ.get("VDA_GOLDEN_ONLY") == "util":
        return
    Offline, Stream = refload.import_reference()
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    torch.set_grad_enabled(False)
    for name, (arch, T, H, W) in FORWARD_CASES.items():
        cfg = ARCH[arch]
        sd = refload.seeded_state_dict(arch, 0)
        m = Offline(encoder=arch, features=cfg["features"], out_channels=cfg["out_channels"]).eval()
        m.load_state_dict(sd, strict=True)
        y = m(window_input(T, H, W)).numpy()
        np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"), depth=y)
        print(name, y.shape, float(y.min()), float(y.max()))
    v = VIDEO_CASE
    cfg = ARCH[v["arch"]]
    sd = refload.seeded_state_dict(v["arch"], 0)
    frames = video_frames(v["n"], v["h"], v["w"])
    m = Offline(encoder=v["arch"], features=cfg["features"], out_channels=cfg["out_channels"]).eval()
    m.load_state_dict(sd, strict=True)
    d, _ = m.infer_video_depth(frames, 30, input_size=v["input_size"], device="cpu", fp32=True)
    mm = Offline(encoder=v["arch"], features=cfg["features"], out_channels=cfg["out_channels"], metric=True).eval()
    mm.load_state_dict(sd, strict=True)
    dm, _ = mm.infer_video_depth(frames, 30, input_size=v["input_size"], device="cpu", fp32=True)
    s = Stream(encoder=v["arch"], features=cfg["features"], out_channels=cfg["out_channels"]).eval()
    s.load_state_dict(sd, strict=True)
    ds = np.stack([s.infer_video_depth_one(frames[i], input_size=v["input_size"], device="cpu", fp32=True)
                   for i in range(v["n"])])
    np.savez_compressed(os.path.join(GOLDEN_DIR, "video_vits_46x44x60.npz"), offline=d, metric=dm, stream=ds)
    print("video", d.shape, dm.shape, ds.shape)


if __name__ == "__main__":
    main()

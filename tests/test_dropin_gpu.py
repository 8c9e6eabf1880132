"""The drop-in contract as the reference's own entry points use it, and BASELINE configs 3-5 at their real sizes.

run.py:45-63 and run_streaming.py:40-100 are restated call for call (checkpoint file -> ctor(**model_configs[enc],
metric=...) -> load_state_dict(torch.load(p, map_location='cpu'), strict=True) -> .to(DEVICE).eval() ->
infer_video_depth / infer_video_depth_one -> save_video / np.savez_compressed); /root/reference does not exist on the
GPU box, so the scripts themselves cannot be executed there (tests/test_abi.py runs their --help in the build
container)."""
import os

import numpy as np
import pytest
import torch

from oracle import make_golden as G
from oracle import refload
from oracle import vda_oracle as O

pytestmark = pytest.mark.gpu
ABSREL_TOL = 1e-2   # BASELINE.json: BF16 path vs the reference FP32 output

MODEL_CONFIGS = {  # run.py:45-49, verbatim
    'vits': {'encoder': 'vits', 'features': 64, 'out_channels': [48, 96, 192, 384]},
    'vitb': {'encoder': 'vitb', 'features': 128, 'out_channels': [96, 192, 384, 768]},
    'vitl': {'encoder': 'vitl', 'features': 256, 'out_channels': [256, 512, 1024, 1024]},
}


def absrel(got, want):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    m = want > 1e-6
    return float(np.mean(np.abs(got[m] - want[m]) / want[m]))


@pytest.fixture(autouse=True)
def _exact_fp32_oracle():
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    yield


@pytest.mark.parametrize("encoder,metric,fp32", [("vits", False, False), ("vits", True, True)])
def test_run_py_call_sequence(tmp_path, encoder, metric, fp32):
    from utils.dc_utils import read_video_frames, save_video
    from video_depth_anything.video_depth import VideoDepthAnything

    DEVICE = 'cuda' if torch.cuda.is_available() else 'cpu'
    sd = refload.seeded_state_dict(encoder, 0)
    ckpt = os.path.join(tmp_path, f"video_depth_anything_{encoder}.pth")
    torch.save(sd, ckpt)
    src = G.video_frames(40, 96, 128, seed=11)
    video = os.path.join(tmp_path, "clip.mp4")
    save_video(src, video, fps=12)                                   # an input file for read_video_frames

    video_depth_anything = VideoDepthAnything(**MODEL_CONFIGS[encoder], metric=metric)
    video_depth_anything.load_state_dict(torch.load(ckpt, map_location='cpu'), strict=True)
    video_depth_anything = video_depth_anything.to(DEVICE).eval()
    frames, target_fps = read_video_frames(video, -1, -1, 1280)
    depths, fps = video_depth_anything.infer_video_depth(frames, target_fps, input_size=112, device=DEVICE, fp32=fp32)

    assert fps == target_fps == 12 and depths.dtype == np.float32 and depths.shape == frames.shape[:3]
    save_video(frames, os.path.join(tmp_path, "clip_src.mp4"), fps=fps)
    save_video(depths, os.path.join(tmp_path, "clip_vis.mp4"), fps=fps, is_depths=True, grayscale=False)
    np.savez_compressed(os.path.join(tmp_path, "clip_depths.npz"), depths=depths)
    assert os.path.getsize(os.path.join(tmp_path, "clip_vis.mp4")) > 0
    with torch.no_grad():
        want = O.infer_video_depth({k: v.cuda() for k, v in sd.items()}, frames, input_size=112, metric=metric,
                                   arch=encoder, device="cuda")
    if fp32:   # north star: FP32 path within rtol 1e-3 of the reference's fp32 forward, every pixel
        assert (np.abs(depths - want) <= 1e-3 * np.abs(want) + 1e-5 * np.abs(want).max()).all()
    else:
        assert absrel(depths, want) <= ABSREL_TOL


def test_run_streaming_py_call_sequence(tmp_path):
    from video_depth_anything.video_depth_stream import VideoDepthAnything

    DEVICE = 'cuda'
    encoder = "vits"
    sd = refload.seeded_state_dict(encoder, 0)
    ckpt = os.path.join(tmp_path, f"video_depth_anything_{encoder}.pth")
    torch.save(sd, ckpt)
    video_depth_anything = VideoDepthAnything(**MODEL_CONFIGS[encoder])
    video_depth_anything.load_state_dict(torch.load(ckpt, map_location='cpu'), strict=True)
    video_depth_anything = video_depth_anything.to(DEVICE).eval()
    frames = G.video_frames(12, 90, 120, seed=5)
    so = O.StreamOracle({k: v.cuda() for k, v in sd.items()}, encoder, device="cuda")
    for i in range(len(frames)):
        depth = video_depth_anything.infer_video_depth_one(frames[i], input_size=112, device=DEVICE, fp32=False)
        assert depth.shape == (90, 120) and depth.dtype == np.float32
        assert absrel(depth, so.step(frames[i], 112)) <= ABSREL_TOL


def test_weight_edit_in_place_invalidates_packed_weights():
    """A parameter changed in place after the first forward must not be served from the stale packed copy."""
    from video_depth_anything.video_depth import VideoDepthAnything

    sd = refload.seeded_state_dict("vits", 0)
    m = VideoDepthAnything(**MODEL_CONFIGS["vits"])
    m.load_state_dict(sd, strict=True)
    m = m.to("cuda").eval()
    x = G.window_input(2, 42, 56).cuda()
    with torch.no_grad():
        a = m(x).clone()
        m.head.scratch.output_conv2[2].bias.add_(0.5)
        b = m(x)
    assert float((b - a).abs().min()) > 0.4


# ------------------------------------------------------------------ BASELINE configs at their real sizes
def test_streaming_518_past_steady_state_vs_oracle():
    """Config 5 at its real frame size: 46 frames of 518x518 through infer_video_depth_one (graph replay from the
    second frame, cache trimming from frame 11) vs the oracle's streaming restatement on the GPU."""
    from video_depth_anything.video_depth_stream import VideoDepthAnything as Stream

    sd = refload.seeded_state_dict("vits", 0)
    m = Stream(**MODEL_CONFIGS["vits"])
    m.load_state_dict(sd, strict=True)
    m = m.to("cuda").eval()
    frames = G.video_frames(46, 518, 518, seed=2)
    so = O.StreamOracle({k: v.cuda() for k, v in sd.items()}, "vits", device="cuda")
    errs = []
    for i in range(len(frames)):
        got = m.infer_video_depth_one(frames[i], input_size=518, device="cuda")
        errs.append(absrel(got, so.step(frames[i], 518)))
    print(f"stream 518x518: absrel first {errs[0]:.2e} max {max(errs):.2e} last {errs[-1]:.2e}")
    assert len(m.frame_cache_list) == 42 and max(errs) <= ABSREL_TOL


def test_vitb_metric_720p_driver_vs_oracle():
    """Config 4 at its real frame size: 1280x720 frames (GPU cubic resize to 518x924, 37x66 tokens, bicubic
    pos-embed resample), metric model (alignment skipped), 24 frames = 2 windows, vs the oracle driver on the GPU."""
    from video_depth_anything.video_depth import VideoDepthAnything

    sd = refload.seeded_state_dict("vitb", 0)
    m = VideoDepthAnything(**MODEL_CONFIGS["vitb"], metric=True)
    m.load_state_dict(sd, strict=True)
    m = m.to("cuda").eval()
    frames = G.video_frames(24, 720, 1280, seed=4)
    got, _ = m.infer_video_depth(frames, 24, input_size=518, device="cuda")
    with torch.no_grad():
        want = O.infer_video_depth({k: v.cuda() for k, v in sd.items()}, frames, input_size=518, metric=True, arch="vitb",
                                   device="cuda")
    assert got.shape == want.shape == (24, 720, 1280)
    e = absrel(got, want)
    print(f"vitb metric 720p driver: absrel {e:.2e}")
    assert e <= ABSREL_TOL

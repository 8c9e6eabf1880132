"""Media I/O with the reference's call signatures (utils/dc_utils.py:18-86), so that run.py and
run_streaming.py import and run unchanged.  decord / imageio / matplotlib are not available here:
frames are read and written with OpenCV, and the depth visualisation (clip-global min/max, 8-bit
quantisation, "inferno" look-up, dc_utils.py:72-86) runs on the GPU (vda_minmax_f32, vda_depth_colorize).
"""
import cv2
import numpy as np


def _scaled_size(h, w, max_res):
    if max_res > 0 and max(h, w) > max_res:
        s = max_res / max(h, w)
        return round(h * s), round(w * s)
    return h, w


def read_video_frames(video_path, process_length, target_fps=-1, max_res=-1):
    """-> (uint8 RGB [N,H,W,3], fps): every `stride`-th frame, at most process_length source frames
    (-1: all), longest side capped at max_res (dc_utils.py:18-67, cv2 branch)."""
    cap = cv2.VideoCapture(video_path)
    if not cap.isOpened():
        raise FileNotFoundError(video_path)
    src_fps = cap.get(cv2.CAP_PROP_FPS)
    h0, w0 = int(cap.get(cv2.CAP_PROP_FRAME_HEIGHT)), int(cap.get(cv2.CAP_PROP_FRAME_WIDTH))
    h, w = _scaled_size(h0, w0, max_res)
    fps = src_fps if target_fps < 0 else target_fps
    stride = max(round(src_fps / fps), 1)
    frames, idx = [], 0
    while True:
        ok, bgr = cap.read()
        if not ok or (process_length > 0 and idx >= process_length):
            break
        if idx % stride == 0:
            rgb = cv2.cvtColor(bgr, cv2.COLOR_BGR2RGB)
            if (h, w) != (h0, w0):
                rgb = cv2.resize(rgb, (w, h))
            frames.append(rgb)
        idx += 1
    cap.release()
    return np.stack(frames, axis=0), fps


def inferno_lut():
    """256 x 3 uint8 RGB table: OpenCV's COLORMAP_INFERNO (pinned by tests/golden/inferno_lut_cv2.npy).
    The reference indexes matplotlib's 256 float "inferno" colours and truncates (`(colormap[q] * 255).astype(uint8)`,
    dc_utils.py:74,79); OpenCV stores the same control points rounded to nearest, so single channels can differ by one
    code value.  matplotlib is not installed on this image (neither here nor on the GPU box), so its float table
    cannot be committed or compared: the index q is bit-exact, the colour is OpenCV's."""
    bgr = cv2.applyColorMap(np.arange(256, dtype=np.uint8).reshape(-1, 1), cv2.COLORMAP_INFERNO)
    return np.ascontiguousarray(bgr[:, 0, ::-1])


def depth_to_vis(depths, grayscale=False):
    """float32 [N,H,W] -> uint8 [N,H,W,3] (or [N,H,W]): clip-global min/max, 8 bits, inferno (GPU kernels)."""
    import torch

    from vda_b200 import lib as L

    L.init()
    d = torch.as_tensor(np.ascontiguousarray(depths, dtype=np.float32)).cuda()
    lut = None if grayscale else torch.from_numpy(inferno_lut()).cuda()
    return L.depth_colorize(d, lut, grey=grayscale).cpu().numpy()


def normalize_depth(depth, near_value, far_value, invert_depth=False):
    """float32 [N,H,W] (or [H,W]) -> float32 same shape: per-frame min/max -> [near, far] remap of the fork's job engine
    (external_engine/engine.py:1099-1141: inverted mapping when far <= near, `invert_depth`, constant frames -> near)
    on the GPU (vda_depth_normalize), bit-exact with the numpy original."""
    import torch

    from vda_b200 import lib as L

    L.init()
    d = np.ascontiguousarray(depth, dtype=np.float32)
    t = torch.from_numpy(d.reshape((-1,) + d.shape[-2:])).cuda()
    return L.depth_normalize(t, near_value, far_value, invert_depth).cpu().numpy().reshape(d.shape)


def save_video(frames, output_video_path, fps=10, is_depths=False, grayscale=False):
    frames = np.asarray(frames)
    vis = depth_to_vis(frames, grayscale) if is_depths else frames
    h, w = vis.shape[1:3]
    writer = cv2.VideoWriter(output_video_path, cv2.VideoWriter_fourcc(*"mp4v"), float(fps), (w, h), isColor=vis.ndim == 4)
    if not writer.isOpened():
        raise OSError(f"cannot open {output_video_path} for writing")
    for f in vis:
        writer.write(cv2.cvtColor(f, cv2.COLOR_RGB2BGR) if f.ndim == 3 else f)
    writer.release()

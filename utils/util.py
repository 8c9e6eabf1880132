"""Import-path parity with the reference's utils/util.py (imported by its video_depth.py:27).
The product path aligns windows on the GPU (vda_b200.parallel.WindowScan); these numpy-facing helpers exist for
callers that import them by name (benchmark/eval-style alignment) and run the same GPU kernels."""
import numpy as np
import torch

from vda_b200 import lib as L


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()


def compute_scale_and_shift(prediction, target, mask=None, scale_only=False):
    """utils/util.py:16-62: least-squares scale (and shift) of `prediction` onto `target` over mask > 0 weights.
    mask None = all ones (the inference path, video_depth.py:138-140).  Sums are accumulated in FP64 on the GPU."""
    L.init()
    p, t = _dev(prediction), _dev(target)
    m = None if mask is None else _dev(mask)
    sums = torch.zeros(5, dtype=torch.float64, device="cuda")
    L.lsq_accum(p, t, p.numel(), sums, mask=m)
    if scale_only:
        s = sums.cpu().numpy()
        return np.float32(s[3] / (s[0] + 1e-6)), 0
    ss = torch.empty(2, device="cuda")
    L.lsq_solve(sums, ss)
    s, t = ss.tolist()
    return np.float32(s), np.float32(t)


def get_interpolate_frames(frame_list_pre, frame_list_post):
    """utils/util.py:65-74 (linear cross-fade, weights i/(n-1)); fp32 roundings as numpy's."""
    assert len(frame_list_pre) == len(frame_list_post)
    n = len(frame_list_pre)
    L.init()
    pre = _dev(np.stack(frame_list_pre))
    post = _dev(np.stack(frame_list_post))
    if not 2 <= n <= 8:
        raise ValueError("get_interpolate_frames: 2..8 frames (INTERP_LEN of video_depth.py:33 is 8)")
    out = torch.empty_like(pre)
    L.blend_frames(pre, post, out, pre[0].numel(), n)
    res = out.cpu().numpy()
    return [res[i] for i in range(n)]

"""Import-path parity with the reference's utils/util.py (imported by its video_depth.py:27).
The product path aligns windows on the GPU (vda_b200.video_depth.align_windows_device); these
numpy-facing helpers exist for callers that import them by name and run the same GPU kernels."""
import numpy as np
import torch

from vda_b200 import lib as L


def compute_scale_and_shift(prediction, target, mask=None, scale_only=False):
    """utils/util.py:16-62 for mask == 1 (the only use on the hot path, video_depth.py:138-140)."""
    if mask is not None and not np.all(mask):
        raise NotImplementedError("only the all-ones mask of the inference path is supported")
    L.init()
    p = torch.from_numpy(np.ascontiguousarray(prediction, dtype=np.float32)).cuda()
    t = torch.from_numpy(np.ascontiguousarray(target, dtype=np.float32)).cuda()
    sums = torch.zeros(5, dtype=torch.float64, device="cuda")
    L.lsq_accum(p, t, p.numel(), sums)
    if scale_only:
        s = sums.cpu().numpy()
        return np.float32(s[3] / (s[0] + 1e-6)), 0
    ss = torch.empty(2, device="cuda")
    L.lsq_solve(sums, ss)
    s, t = ss.tolist()
    return np.float32(s), np.float32(t)


def get_interpolate_frames(frame_list_pre, frame_list_post):
    """utils/util.py:65-74 (linear cross-fade, weights i/(n-1))."""
    assert len(frame_list_pre) == len(frame_list_post)
    n = len(frame_list_pre)
    L.init()
    pre = torch.from_numpy(np.stack(frame_list_pre).astype(np.float32)).cuda()
    post = torch.from_numpy(np.stack(frame_list_post).astype(np.float32)).cuda()
    ident = torch.tensor([1.0, 0.0], device="cuda")
    # blend kernel clamps post at 0 after the (identity) affine; the reference's callers pass
    # already-clamped frames, for which this is the identity.
    L.affine_clamp_blend(pre, post, pre[0].numel(), n, ident)
    out = pre.cpu().numpy()
    return [out[i] for i in range(n)]

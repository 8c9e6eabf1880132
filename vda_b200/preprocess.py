"""Frame pre-processing (util/transform.py:51-157 as composed in video_depth.py:77-89):
keep-aspect "lower_bound" resize to a multiple of 14 with cv2 INTER_CUBIC on float32 [0,1] RGB,
ImageNet normalisation evaluated in float64 then cast to float32, HWC -> CHW.

The reference re-runs this for every slot of every window on one Python thread (32 cv2 resizes per
window: 10x slower than the GPU forward at 1280x720).  Here every source frame is converted exactly
once: the uint8 frames go to the device as they are (a quarter of the fp32 bytes) and
`vda_preprocess_u8` (csrc/preprocess.cu) does resize + normalise + transpose in one kernel, restating
OpenCV's float32 cubic arithmetic.  `preprocess_frames_cv2` is the reference's own host path, kept as
the comparand for tests.
"""
from concurrent.futures import ThreadPoolExecutor
import os

import cv2
import numpy as np
import torch

_MEAN = np.array([0.485, 0.456, 0.406])
_STD = np.array([0.229, 0.224, 0.225])
_POOL = None


def net_input_size(h0, w0, input_size=518):
    """Network input (height, width) for a source frame (video_depth.py:71-75, transform.py:51-107)."""
    ratio = max(h0, w0) / min(h0, w0)
    if ratio > 1.78:
        input_size = int(input_size * 1.777 / ratio)
        input_size = round(input_size / 14) * 14
    scale = max(input_size / w0, input_size / h0)

    def snap(v):
        y = int(np.round(v / 14) * 14)
        return y if y >= input_size else int(np.ceil(v / 14) * 14)

    return snap(scale * h0), snap(scale * w0)


def _one(frame, nh, nw, dst):
    img = frame.astype(np.float32) / 255.0
    img = cv2.resize(img, (nw, nh), interpolation=cv2.INTER_CUBIC)
    img = (img - _MEAN) / _STD
    dst[...] = np.transpose(img, (2, 0, 1))


def preprocess_frames(frames, indices, nh, nw, device):
    """frames uint8 [N,H0,W0,3] (host); returns f32 [len(indices),3,nh,nw] on `device` (GPU kernel)."""
    from . import lib as L

    frames = np.asarray(frames)
    sel = torch.from_numpy(np.ascontiguousarray(frames[list(indices)]))
    host = torch.empty(sel.shape, dtype=torch.uint8, pin_memory=True)
    host.copy_(sel)
    dev_u8 = host.to(device, non_blocking=True)
    out = torch.empty(len(indices), 3, nh, nw, dtype=torch.float32, device=device)
    L.preprocess_u8(dev_u8, out, nh, nw)
    return out


_FEED_POOL = None


def _feed_pool():
    global _FEED_POOL
    if _FEED_POOL is None:
        _FEED_POOL = ThreadPoolExecutor(max_workers=FrameFeeder.THREADS)
    return _FEED_POOL


class FrameFeeder:
    """Host -> device feed of the window loop (video_depth.py:92-107): fills the network input of one window.

    The window's source frames are gathered as uint8 into one of two pinned staging slots (allocated once per
    frame size and reused: pinning ~1 GB per call costs more than a whole second on the GPU box), copied to the
    device on a side stream while the previous window still computes, and converted by `vda_preprocess_u8` straight
    into the window's input tensor.  When the previous window of the video ran on this GPU too, its key-frame slots
    ARE this window's first ten slots (video_depth.py:96-99: `pre_input[:, KEYFRAMES]`), so only the 22 new frames
    are uploaded.  Nothing of size O(video) stays on the GPU."""

    SLOT_FRAMES = 32
    THREADS = 4

    def __init__(self, h0, w0, nh, nw, device):
        self.key = (h0, w0, nh, nw, str(device))
        self.device = device
        self.nh, self.nw = nh, nw
        self.pinned = [torch.empty(self.SLOT_FRAMES, h0, w0, 3, dtype=torch.uint8, pin_memory=True) for _ in range(2)]
        self.staged = [torch.empty(self.SLOT_FRAMES, h0, w0, 3, dtype=torch.uint8, device=device) for _ in range(2)]
        self.copied = [None, None]     # H2D of the slot has finished: the pinned half may be rewritten
        self.consumed = [None, None]   # the pre-process kernel has read the device half
        self.turn = 0
        self.side = torch.cuda.Stream(device)
        self.key_index = None

    def fill(self, frames, indices, x, key_slots=None):
        """x f32 [len(indices),3,nh,nw] <- network inputs of frames[indices].  key_slots: if given, x still holds the
        previous window's input and x[key_slots] become slots 0..len(key_slots)-1 (the caller guarantees that these
        are the same source frames); only the remaining slots are uploaded.  Asynchronous on the current stream."""
        from . import lib as L

        first = 0
        if key_slots is not None:
            if self.key_index is None or self.key_index.numel() != len(key_slots):
                self.key_index = torch.tensor(list(key_slots), dtype=torch.long, device=self.device)
            first = len(key_slots)
            x[:first].copy_(x.index_select(0, self.key_index))
        part = list(indices[first:])
        m = len(part)
        if m == 0:
            return
        t = self.turn
        self.turn ^= 1
        if self.copied[t] is not None:
            self.copied[t].synchronize()
        dst = self.pinned[t].numpy()
        # Host gather into the pinned slot on a few threads (numpy releases the GIL while copying).  One np.take of a
        # window's 32 frames (26 MB at 518 x 518) took 4-17 ms with eight ranks on one host; it sits on the critical
        # path of a video's first round and skews the ranks of a lock-step round (8-GPU trace).
        runs, a0 = [], 0
        for i in range(1, m + 1):          # maximal runs of consecutive source frames -> block copies
            if i == m or part[i] != part[i - 1] + 1:
                runs.append((a0, i))
                a0 = i
        jobs = []
        for a, b in runs:
            step = max(1, (b - a + self.THREADS - 1) // self.THREADS) if len(runs) < self.THREADS else b - a
            jobs += [(c, min(c + step, b)) for c in range(a, b, step)]

        def _copy(job):
            a, b = job
            dst[a:b] = frames[part[a]:part[a] + (b - a)]

        if m >= 8 and len(jobs) > 1:
            list(_feed_pool().map(_copy, jobs))
        else:
            for job in jobs:
                _copy(job)
        main = torch.cuda.current_stream()
        with torch.cuda.stream(self.side):
            if self.consumed[t] is not None:
                self.side.wait_event(self.consumed[t])
            self.staged[t][:m].copy_(self.pinned[t][:m], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
        self.copied[t] = ev
        main.wait_event(ev)
        L.preprocess_u8(self.staged[t][:m], x[first:first + m], self.nh, self.nw)
        done = torch.cuda.Event()
        done.record()
        self.consumed[t] = done


def preprocess_frames_cv2(frames, indices, nh, nw, device):
    """The reference's host path (cv2 on a thread pool); test comparand of preprocess_frames."""
    global _POOL
    if _POOL is None:
        _POOL = ThreadPoolExecutor(max_workers=min(16, os.cpu_count() or 1))
    host = torch.empty(len(indices), 3, nh, nw, dtype=torch.float32, pin_memory=torch.cuda.is_available())
    arr = host.numpy()
    list(_POOL.map(lambda j: _one(frames[indices[j]], nh, nw, arr[j]), range(len(indices))))
    return host.to(device, non_blocking=True)

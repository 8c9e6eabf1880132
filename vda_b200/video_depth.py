"""Drop-in `VideoDepthAnything` (offline, 32-frame windows) on the B200 kernel library.

Mirrors the reference's public surface (video_depth_anything/video_depth.py:35-162): same ctor
kwargs, same state_dict keys, `forward(x[B,T,3,H,W]) -> [B,T,H,W]`, and
`infer_video_depth(frames, target_fps, input_size, device, fp32) -> (float32 [N,H0,W0], fps)`.
What differs is underneath: every window is a fixed launch sequence of libvda_b200 kernels
(vda_b200.engine); each source frame is pre-processed once; all windows' depth maps stay on the
GPU; the sequential scale/shift alignment and the cross-fade run as GPU kernels without host
synchronisation; one device->host copy returns the result.
"""
import os
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch
from torch import nn

from . import lib as L
from . import parallel
from .engine import Engine
from .modules import ARCH, DinoEncoder, DPTTemporalHead
from .preprocess import FrameFeeder, net_input_size, preprocess_frames

# window semantics of the reference, "do not change" (video_depth.py:29-33)
INFER_LEN = 32
OVERLAP = 10
KEYFRAMES = [0, 12, 24, 25, 26, 27, 28, 29, 30, 31]
INTERP_LEN = 8

# ctor kwargs per encoder, as every caller of the reference spells them (run.py:45-49)
MODEL_CONFIGS = {
    "vits": {"encoder": "vits", "features": 64, "out_channels": [48, 96, 192, 384]},
    "vitb": {"encoder": "vitb", "features": 128, "out_channels": [96, 192, 384, 768]},
    "vitl": {"encoder": "vitl", "features": 256, "out_channels": [256, 512, 1024, 1024]},
}


def randomize_zero_init_(model, seed=0):
    """Random-init benchmarking helper: the reference zero-initialises every motion module's
    proj_out (motion_module.py:57-58), which would make the temporal branch a no-op on fresh weights;
    give it small random values so a random-init model exercises the whole path."""
    g = torch.Generator().manual_seed(seed)
    for m in model.head.motion_modules:
        po = m.temporal_transformer.proj_out
        with torch.no_grad():
            po.weight.copy_(torch.randn(po.weight.shape, generator=g) * po.weight.shape[1] ** -0.5)
            po.bias.copy_(torch.randn(po.bias.shape, generator=g) * 0.02)
    # keep the regressed depth strictly positive on random weights (the last 1x1 conv sums ReLU outputs:
    # non-negative weights + a positive bias cannot be zeroed by the final ReLU)
    last = model.head.scratch.output_conv2[2]
    with torch.no_grad():
        last.weight.abs_()
        last.bias.fill_(0.1)
    return model


def window_frame_indices(n_frames):
    """Raw-frame indices of every window (closed form of video_depth.py:92-115: the keyframe slots
    are copied from the previous window's already-substituted input, so slot 0 is always frame 0)."""
    step = INFER_LEN - OVERLAP
    last = n_frames - 1
    wins = []
    for start in range(0, n_frames, step):
        cur = [min(start + i, last) for i in range(INFER_LEN)]
        if wins:
            cur[:OVERLAP] = [wins[-1][j] for j in KEYFRAMES]
        wins.append(cur)
    return wins


class VideoDepthAnything(nn.Module):
    def __init__(self, encoder="vitl", features=256, out_channels=[256, 512, 1024, 1024], use_bn=False,
                 use_clstoken=False, num_frames=32, pe="ape", metric=False):
        super().__init__()
        if use_bn or use_clstoken or pe != "ape":
            raise NotImplementedError("use_bn / use_clstoken / pe='rope' are never selected by the reference's "
                                      "callers and are out of scope (SURVEY.md section 2)")
        if encoder not in ARCH:
            raise KeyError(encoder)
        self.intermediate_layer_idx = {k: list(v["taps"]) for k, v in ARCH.items()}
        self.encoder = encoder
        self.pretrained = DinoEncoder(encoder)
        self.head = DPTTemporalHead(self.pretrained.embed_dim, features, out_channels, num_frames)
        self.metric = metric
        self._engine = None
        self._engine32 = None
        # forward() precision: "bf16" (autocast path of the reference, video_depth.py:108) or "fp32"
        # (the reference's forward outside autocast); infer_video_depth(fp32=...) selects it per call.
        self.precision = "bf16"

    # -- lifecycle: packed weights are rebuilt lazily after .to()/.cuda()/load_state_dict()
    def _apply(self, fn, *a, **k):
        self._engine = self._engine32 = None
        self._tensors = None
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, *a, **k):
        self._engine = self._engine32 = None
        return super().load_state_dict(*a, **k)

    def _weights_version(self):
        """Changes whenever a parameter or buffer is modified in place through autograd-visible ops
        (copy_, fill_, optimizer steps, sub-module load_state_dict): the packed kernel-layout weights and the
        captured CUDA graphs are then rebuilt instead of silently replaying stale weights."""
        ts = getattr(self, "_tensors", None)
        if ts is None:
            ts = self._tensors = list(self.parameters()) + list(self.buffers())
        return sum(t._version for t in ts)

    def engine(self, fp32=False):
        """Window executor for the module's current device; fp32=True -> the FP32-accuracy executor
        (split-bf16 tensor-core products, fp32 everywhere else; engine.Engine)."""
        dev = self.pretrained.cls_token.device
        v = self._weights_version()
        if v != getattr(self, "_packed_version", None):
            self._engine = self._engine32 = None
            self._packed_version = v
        if fp32:
            if self._engine32 is None or self._engine32.device != dev:
                self._engine32 = Engine(self, dev, fp32=True)
            return self._engine32
        if self._engine is None or self._engine.device != dev:
            self._engine = Engine(self, dev)
        return self._engine

    def forward(self, x, fp32=None):
        """x f32 [B,T,3,H,W], H and W multiples of 14 -> depth f32 [B,T,H,W] (>= 0).
        fp32=None follows self.precision."""
        B, T = x.shape[:2]
        eng = self.engine(self.precision == "fp32" if fp32 is None else fp32)
        out = [eng.forward(x[b].float()).clone() for b in range(B)]
        return torch.stack(out, 0)

    @torch.no_grad()
    def forward_pipelined(self, host_windows, out=None, fp32=False):
        """Window forwards with HOST buffers in and out, copies overlapped with compute.
        host_windows: iterable of pinned f32 [T,3,H,W] tensors; out: optional list of pinned f32 [T,H,W]
        tensors to fill (allocated otherwise).  Window i+1's H2D copy and window i-1's D2H copy run on
        two side streams while window i computes; two staging buffers each way.  Returns the list of
        host depth tensors (valid after the call: it synchronises at the end)."""
        eng = self.engine(fp32)
        dev = eng.device
        outs = []
        with torch.cuda.device(dev):
            main = torch.cuda.current_stream()
            s_in, s_out = self._side_streams(dev)
            st = None
            for i, xh in enumerate(host_windows):
                if st is None or st["shape"] != tuple(xh.shape):
                    T, _, H, W = xh.shape
                    st = dict(shape=tuple(xh.shape),
                              xin=[torch.empty(T, 3, H, W, device=dev) for _ in range(2)],
                              dout=[torch.empty(T, H, W, device=dev) for _ in range(2)],
                              in_ready=[torch.cuda.Event() for _ in range(2)], in_free=[None, None],
                              out_ready=[torch.cuda.Event() for _ in range(2)], out_free=[None, None])
                b = i & 1
                oh = out[i] if out is not None else torch.empty(st["shape"][0], *st["shape"][2:], pin_memory=True)
                with torch.cuda.stream(s_in):
                    if st["in_free"][b] is not None:
                        s_in.wait_event(st["in_free"][b])       # window i-2 has consumed this staging buffer
                    st["xin"][b].copy_(xh, non_blocking=True)
                    st["in_ready"][b].record(s_in)
                main.wait_event(st["in_ready"][b])
                d = eng.forward(st["xin"][b])
                st["in_free"][b] = torch.cuda.Event()
                st["in_free"][b].record(main)
                if st["out_free"][b] is not None:
                    main.wait_event(st["out_free"][b])          # window i-2's D2H has drained this buffer
                st["dout"][b].copy_(d)                          # d is a workspace buffer the next window overwrites
                st["out_ready"][b].record(main)
                with torch.cuda.stream(s_out):
                    s_out.wait_event(st["out_ready"][b])
                    oh.copy_(st["dout"][b], non_blocking=True)
                    st["out_free"][b] = torch.cuda.Event()
                    st["out_free"][b].record(s_out)
                outs.append(oh)
            s_out.synchronize()
            main.synchronize()
        return outs

    def _side_streams(self, dev):
        key = str(dev)
        if getattr(self, "_streams", None) is None or self._streams[0] != key:
            self._streams = (key, torch.cuda.Stream(dev), torch.cuda.Stream(dev))
        return self._streams[1], self._streams[2]

    # ------------------------------------------------------------------ offline driver
    @torch.no_grad()
    def infer_video_depth(self, frames, target_fps, input_size=518, device="cuda", fp32=False, original_width=None,
                          original_height=None):
        """frames uint8 [N,H0,W0,3] RGB -> (float32 [N,H0,W0], target_fps).

        fp32=False: BF16 tensor-core path (the reference's autocast branch, video_depth.py:108).
        fp32=True: FP32-accuracy path (rtol 1e-3 against the reference's fp32 forward).
        original_width / original_height (both given): the aligned maps are upscaled bilinearly
        (align_corners=True) to that size before they leave the GPU -- the kwargs of the fork's engine copy,
        external_engine/video_depth_anything/video_depth.py:81-90, :181-203.

        One pass, pipelined per window: upload + pre-process the window's frames, forward, resize to the source
        size, one step of the alignment scan (vda_b200.parallel.WindowScan), then the frames that step made final go
        device -> pinned ring -> the result array on a side stream while the next window computes.

        Under torch.distributed (one process per GPU) the windows are dealt round-robin over the ranks and the result
        goes where vda_b200.parallel.set_result_rank / set_result_sharded say: every rank (default), one rank (the
        others return (None, fps)), or each rank only the frames of its own windows (the rest of the returned array
        is uninitialised; `self.last_owned_frames` lists the [lo, hi) ranges that are valid)."""
        frames = np.asarray(frames)
        n, h0, w0 = frames.shape[:3]
        up = (int(original_height), int(original_width)) if original_width is not None and original_height is not None else None
        dev = self._device(device)
        world, _ = parallel.world_rank()
        want = parallel.wants_result()
        out = np.empty((n,) + (up or (h0, w0)), np.float32) if want else None
        dl = None
        owned = []
        if want:
            # slots for two rounds of final frames: one round being drained on the host while the next is enqueued
            per_round = world if (world > 1 and parallel.result_mode() != "sharded") else 1
            dl = self._downloader(dev, h0, w0, up, 2 * per_round)
            dl.begin(out)

        def sink(lo, maps):
            m = min(maps.shape[0], n - lo)
            if m > 0:
                dl.push(maps[:m], lo)
                owned.append((lo, lo + m))

        # the scan owns the device buffers the downloader's side-stream copies read: keep it until they have drained
        scan = self._run_video(frames, input_size, dev, fp32, sink if want else None, dl.next_round if want else None)
        if want:
            dl.finish()
        del scan
        self.last_owned_frames = owned
        return out, target_fps

    @torch.no_grad()
    def infer_video_depth_device(self, frames, input_size=518, device="cuda", fp32=False):
        """Same as infer_video_depth but leaves the aligned depth [N,H0,W0] on the GPU (None on ranks that do not
        receive the result)."""
        frames = np.asarray(frames)
        n, h0, w0 = frames.shape[:3]
        dev = self._device(device)
        want = parallel.wants_result()
        out = torch.empty(n, h0, w0, dtype=torch.float32, device=dev) if want else None

        def sink(lo, maps):
            m = min(maps.shape[0], n - lo)
            if m > 0:
                out[lo:lo + m].copy_(maps[:m])

        self._run_video(frames, input_size, dev, fp32, sink if want else None, None)
        return out

    def _device(self, device):
        dev = torch.device(device if device != "cuda" else f"cuda:{torch.cuda.current_device()}")
        if self.pretrained.cls_token.device != dev:
            raise RuntimeError(f"model is on {self.pretrained.cls_token.device}, call .to('{dev}') first")
        return dev

    def _run_video(self, frames, input_size, dev, fp32, sink, end_of_round):
        """The window loop (video_depth.py:92-160).  Round j: this rank's window k = j*world + rank is fed,
        forwarded, resized to the source size and pushed into the alignment scan together with the other ranks'
        windows of the round; sink(first_frame, maps) receives the frames that became final; end_of_round() is
        called after a round's GPU work has been enqueued (the downloader drains the PREVIOUS round there, so the
        host-side copies of round j-1 run while the GPU computes round j)."""
        n, h0, w0 = frames.shape[:3]
        nh, nw = net_input_size(h0, w0, input_size)
        wins = window_frame_indices(n)
        world, rank = parallel.world_rank()
        eng = self.engine(fp32)
        timing = self.last_timing = {"windows": len(wins), "rounds": parallel.n_rounds(len(wins), world), "world": world}
        with torch.cuda.device(dev):
            main = torch.cuda.current_stream()
            main.synchronize()
            t0 = time.perf_counter()
            e_begin, e_last = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e_begin.record()
            feeder = getattr(self, "_feeder", None)
            if feeder is None or feeder.key != (h0, w0, nh, nw, str(dev)):
                feeder = self._feeder = FrameFeeder(h0, w0, nh, nw, dev)
            key = ("video", h0, w0, nh, nw, str(dev))
            bufs = getattr(self, "_video_bufs", None)
            if bufs is None or bufs[0] != key:
                bufs = self._video_bufs = (key, torch.empty(INFER_LEN, 3, nh, nw, dtype=torch.float32, device=dev),
                                           [torch.empty(INFER_LEN, h0, w0, dtype=torch.float32, device=dev) for _ in range(2)])
            x, raws = bufs[1], bufs[2]
            scan = parallel.WindowScan(len(wins), h0, w0, GpuAlignOps(dev), self.metric)
            prev_k = None
            trace = [] if os.environ.get("VDA_VIDEO_TRACE") == "1" else None   # per-round host / device times (ms)
            for j in range(timing["rounds"]):
                k = j * world + rank
                raw = None
                h = [time.perf_counter()] if trace is not None else None
                ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if trace is not None else None
                if trace is not None:
                    ev[0].record()
                if k < len(wins):
                    feeder.fill(frames, wins[k], x, KEYFRAMES if prev_k is not None and prev_k == k - 1 else None)
                    prev_k = k
                    if trace is not None:
                        h.append(time.perf_counter())
                        ev[1].record()
                    d = eng.forward(x)
                    raw = raws[j & 1]
                    L.bilinear_f32(d, raw, INFER_LEN, nh, nw, h0, w0)
                    e_last = torch.cuda.Event(enable_timing=True)
                    e_last.record()
                elif trace is not None:
                    h.append(time.perf_counter())
                    ev[1].record()
                if trace is not None:
                    h.append(time.perf_counter())
                    ev[2].record()
                for lo, maps in scan.push(j, raw):
                    if sink is not None:
                        sink(lo, maps)
                if trace is not None:
                    h.append(time.perf_counter())
                    ev[3].record()
                if end_of_round is not None:
                    end_of_round()
                if trace is not None:
                    h.append(time.perf_counter())
                    trace.append((h, ev))
            main.synchronize()
            timing["seconds_device"] = time.perf_counter() - t0
            timing["seconds_to_last_forward"] = e_begin.elapsed_time(e_last) * 1e-3
            if trace is not None:
                timing["trace"] = [{"host_ms": {"fill": (h[1] - h[0]) * 1e3, "forward_enqueue": (h[2] - h[1]) * 1e3,
                                                "push_enqueue": (h[3] - h[2]) * 1e3, "end_of_round": (h[4] - h[3]) * 1e3},
                                    "device_ms": {"since_begin": e_begin.elapsed_time(ev[0]), "fill": ev[0].elapsed_time(ev[1]),
                                                  "forward": ev[1].elapsed_time(ev[2]), "push": ev[2].elapsed_time(ev[3])}}
                                   for h, ev in trace]
        return scan

    def _downloader(self, dev, h0, w0, up, slots):
        key = (str(dev), h0, w0, up, slots)
        dl = getattr(self, "_dl", None)
        if dl is None or dl.key != key:
            dl = self._dl = HostDownloader(dev, h0, w0, up, slots)
        return dl


class GpuAlignOps:
    """The arithmetic of vda_b200.parallel.WindowScan on the CUDA kernels of csrc/align.cu."""

    def __init__(self, device):
        self.device = device
        self.sums = torch.zeros(5, dtype=torch.float64, device=device)

    def empty(self, *shape):
        return torch.empty(*shape, dtype=torch.float32, device=self.device)

    def scale_shift_table(self, n):
        t = torch.zeros(n, 2, dtype=torch.float32, device=self.device)
        t[:, 0] = 1.0
        return t

    def scale_shift(self, pred, target, ss, metric):
        """ss <- least-squares (scale, shift) of pred onto target (utils/util.py:40-62), FP64 sums; the solve leaves
        the sums zeroed for the next window."""
        L.lsq_accum(pred, target, pred.numel(), self.sums)
        L.lsq_solve(self.sums, ss, metric, reset=True)

    def affine_clamp(self, src, dst, ss):
        L.affine_clamp(src, dst, src.numel(), ss)

    def affine_clamp_blend(self, pre, post, ss):
        L.affine_clamp_blend(pre, post, pre[0].numel(), pre.shape[0], ss)


def align_windows_device(raw, metric=False):
    """All windows at once on one GPU: raw f32 [n_windows*32, H0, W0] -> f32 [32 + 22*(n_windows-1), H0, W0]."""
    n_win = raw.shape[0] // INFER_LEN
    scan = parallel.WindowScan(n_win, raw.shape[1], raw.shape[2], GpuAlignOps(raw.device), metric, group=parallel.SOLO)
    out = torch.empty(INFER_LEN + (INFER_LEN - OVERLAP) * (n_win - 1), raw.shape[1], raw.shape[2], dtype=torch.float32,
                      device=raw.device)
    for k in range(n_win):
        for lo, maps in scan.push(k, raw[k * INFER_LEN:(k + 1) * INFER_LEN]):
            out[lo:lo + maps.shape[0]].copy_(maps)
    return out


_POOL = None


def _copy_pool():
    global _POOL
    if _POOL is None:
        _POOL = ThreadPoolExecutor(max_workers=HostDownloader.MAX_THREADS)
    return _POOL


class HostDownloader:
    """Device -> host path of the result (video_depth.py:113,162): batches of <= 32 final maps go through a ring
    of pinned slots on a side stream and are copied into the caller's (pageable) array one round later, while the
    GPU works on the next window.  The ring is allocated once per frame size and reused across calls (pinning is
    slow: ~1 s per GB on the GPU box); the optional bilinear upscale to the caller's plate size runs on the main
    stream before the copy."""

    FRAMES = INFER_LEN
    THREADS = 4          # host copy threads per drain; MAX_THREADS when this rank alone collects a multi-rank result
    MAX_THREADS = 8

    def __init__(self, dev, h0, w0, up=None, slots=2):
        self.key = (str(dev), h0, w0, up, slots)
        self.dev, self.h0, self.w0, self.up = dev, h0, w0, up
        oh, ow = up or (h0, w0)
        self.ring = [torch.empty(self.FRAMES, oh, ow, dtype=torch.float32, pin_memory=True) for _ in range(slots)]
        self.upbuf = [torch.empty(self.FRAMES, oh, ow, dtype=torch.float32, device=dev) for _ in range(slots)] if up else None
        self.side = torch.cuda.Stream(dev)
        self.begin(None)

    def begin(self, out):
        self.out = out
        # one rank collecting every rank's frames drains world x more bytes per round while the other ranks' hosts idle
        self.threads = self.MAX_THREADS if (parallel.active() and isinstance(parallel.result_mode(), int)) else self.THREADS
        self.pending = [None] * len(self.ring)
        self.turn = 0                         # pushes so far; push i uses slot i % len(ring)
        self.prev_start = self.round_start = 0  # pushes [round_start, turn) belong to the round being enqueued

    def _drain(self, t):
        if self.pending[t] is not None:
            ev, lo, m = self.pending[t]
            ev.synchronize()
            src = self.ring[t].numpy()
            if m >= 8:      # numpy releases the GIL while copying: a few threads triple the host copy rate
                cuts = [m * i // self.threads for i in range(self.threads + 1)]
                list(_copy_pool().map(lambda i: self.out[lo + cuts[i]:lo + cuts[i + 1]].__setitem__(
                    slice(None), src[cuts[i]:cuts[i + 1]]), range(self.threads)))
            else:
                self.out[lo:lo + m] = src[:m]
            self.pending[t] = None

    def push(self, src, lo):
        """src: device f32 [m <= 32, h0, w0], final; becomes out[lo:lo+m]."""
        t = self.turn % len(self.ring)
        self.turn += 1
        self._drain(t)
        m = src.shape[0]
        with torch.cuda.device(self.dev):
            if self.up:
                L.bilinear_f32(src.contiguous(), self.upbuf[t][:m], m, self.h0, self.w0, self.up[0], self.up[1])
                src = self.upbuf[t][:m]
            ready = torch.cuda.Event()
            ready.record()
            with torch.cuda.stream(self.side):
                self.side.wait_event(ready)
                self.ring[t][:m].copy_(src, non_blocking=True)
                done = torch.cuda.Event()
                done.record()
        self.pending[t] = (done, lo, m)

    def next_round(self):
        """Called when a round's GPU work has been enqueued: copies the PREVIOUS round's maps into the result
        array (their device -> pinned copies finished while this round was being enqueued or will shortly)."""
        for i in range(max(self.prev_start, self.turn - len(self.ring)), self.round_start):
            self._drain(i % len(self.ring))
        self.prev_start, self.round_start = self.round_start, self.turn

    def finish(self):
        for t in range(len(self.ring)):
            self._drain((self.turn + t) % len(self.ring))
        self.out = None

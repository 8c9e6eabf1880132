"""Drop-in streaming `VideoDepthAnything` (video_depth_anything/video_depth_stream.py:31-161):
frame-by-frame inference with a temporal cache.  Same ctor (no `metric` kwarg, :33-42), same
state_dict keys, same stateful attributes (`transform`, `frame_id_list`, `frame_cache_list`, `gap`,
`id`, `frame_height`, `frame_width`) and the same list bookkeeping; the cache entries are slot
numbers into a device-resident K/V cache (vda_b200.engine.StreamCache) instead of lists of tensors.
"""
import numpy as np
import torch
from torch import nn

from . import lib as L
from .engine import Engine, StreamCache
from .modules import ARCH, DinoEncoder, DPTTemporalHead
from .preprocess import net_input_size
from .video_depth import INFER_LEN, INTERP_LEN, OVERLAP


class VideoDepthAnything(nn.Module):
    def __init__(self, encoder="vitl", features=256, out_channels=[256, 512, 1024, 1024], use_bn=False,
                 use_clstoken=False, num_frames=32, pe="ape"):
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
        self.transform = None
        self.frame_id_list = []
        self.frame_cache_list = []
        self.gap = (INFER_LEN - OVERLAP) * 2 - 1 - (OVERLAP - INTERP_LEN)
        assert self.gap == 41
        self.id = -1
        self._engine = None
        self._cache = None
        self._io = None

    def _apply(self, fn, *a, **k):
        self._engine = None
        self._tensors = None
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, *a, **k):
        self._engine = None
        return super().load_state_dict(*a, **k)

    def engine(self, fp32=False):
        """One executor per precision (BF16 tensor-core path / FP32-accuracy path), built on first use and
        rebuilt when a parameter was modified in place since (see video_depth.VideoDepthAnything._weights_version)."""
        dev = self.pretrained.cls_token.device
        ts = getattr(self, "_tensors", None)
        if ts is None:
            ts = self._tensors = list(self.parameters()) + list(self.buffers())
        v = sum(t._version for t in ts)
        if v != getattr(self, "_packed_version", None):
            self._engine = None
            self._packed_version = v
        if self._engine is None or self._engine.get("dev") != dev:
            self._engine = {"dev": dev}
        if fp32 not in self._engine:
            self._engine[fp32] = Engine(self, dev, fp32=fp32)
        return self._engine[fp32]

    def forward(self, x):
        """Cache-free forward, x f32 [B,T,3,H,W] -> [B,T,H,W] (video_depth_stream.py:62-63)."""
        eng = self.engine()
        return torch.stack([eng.forward(x[b].float()).clone() for b in range(x.shape[0])], 0)

    def _free_slot(self):
        used = set(self.frame_cache_list)
        for s in range(StreamCache.CAP):
            if s not in used:
                return s
        raise RuntimeError("streaming cache exhausted")

    @torch.no_grad()
    def infer_video_depth_one(self, frame, input_size=518, device="cuda", fp32=False):
        """frame uint8 [H0,W0,3] RGB -> float32 [H0,W0] depth of that frame (video_depth_stream.py:77-161)."""
        self.id += 1
        dev = self.pretrained.cls_token.device
        frame = np.asarray(frame)
        if self.transform is None:  # first frame
            self.frame_height, self.frame_width = frame.shape[:2]
            self.transform = net_input_size(self.frame_height, self.frame_width, input_size)
        else:
            assert frame.shape[0] == self.frame_height
            assert frame.shape[1] == self.frame_width
        nh, nw = self.transform
        eng = self.engine(bool(fp32))
        with torch.cuda.device(dev):
            if self._cache is None or self.id == 0:
                self._cache = StreamCache(eng, nh // 14, nw // 14)
            elif self._cache.fp32 != eng.fp32:
                raise RuntimeError("infer_video_depth_one: `fp32` must stay the same for all frames of a stream "
                                   "(the temporal cache holds projections in that precision)")
            io = self._io
            if io is None or io["key"] != (self.frame_height, self.frame_width, nh, nw, str(dev)):
                # per-stream staging, allocated once: pinned frame in, device frame, network input, resized depth,
                # pinned depth out (a pinned allocation per frame costs more than the pre-processing kernel)
                h0, w0 = self.frame_height, self.frame_width
                io = self._io = dict(
                    key=(h0, w0, nh, nw, str(dev)),
                    pin_in=torch.empty(1, h0, w0, 3, dtype=torch.uint8, pin_memory=True),
                    dev_in=torch.empty(1, h0, w0, 3, dtype=torch.uint8, device=dev),
                    x=torch.empty(1, 3, nh, nw, dtype=torch.float32, device=dev),
                    out=torch.empty(1, h0, w0, dtype=torch.float32, device=dev),
                    pin_out=torch.empty(h0, w0, dtype=torch.float32, pin_memory=True))
            io["pin_in"].numpy()[0] = frame
            io["dev_in"].copy_(io["pin_in"], non_blocking=True)
            L.preprocess_u8(io["dev_in"], io["x"], nh, nw)
            x = io["x"]
            slot = self._free_slot()
            if self.id == 0:
                cur = []
            else:
                cur = self.frame_cache_list[0:2] + self.frame_cache_list[-INFER_LEN + 3:]
                assert len(cur) == INFER_LEN - 1
            self._cache.begin_frame(cur, slot)
            d = eng.forward(x, stream_ctx=self._cache)
            L.bilinear_f32(d, io["out"], 1, nh, nw, self.frame_height, self.frame_width)
            io["pin_out"].copy_(io["out"][0], non_blocking=True)
            torch.cuda.current_stream().synchronize()
            new_depth = io["pin_out"].numpy().copy()
        if self.id == 0:
            # copy the cache to simulate a full window (video_depth_stream.py:117-119)
            self.frame_cache_list = [slot] * INFER_LEN
            self.frame_id_list.extend([0] * (INFER_LEN - 1))
        else:
            self.frame_cache_list.append(slot)
        # adjust the sliding window (video_depth_stream.py:155-159)
        self.frame_id_list.append(self.id)
        if self.id + INFER_LEN > self.gap + 1:
            del self.frame_id_list[1]
            del self.frame_cache_list[1]
        return new_depth

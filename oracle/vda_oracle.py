"""ORACLE -- test infrastructure only.  Never imported by the product path (vda_b200/,
video_depth_anything/, utils/); only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may use it.

A functional restatement, in plain PyTorch fp32 ops, of the reference's per-window inference path
(apanner/VDA, paths relative to the reference root).  It works directly from a state_dict with the
reference's key names, so it has no module classes of its own.  Parity pinning: the reference has
no tests/golden vectors for this path; the pin is the reference module itself,
run in the build container by oracle/make_golden.py on seeded weights/inputs, whose outputs are
committed under tests/golden/ and compared by tests/test_oracle.py.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

ARCH = {
    # video_depth.py:49-53, run.py:45-49, dinov2.py:339-378
    "vits": dict(dim=384, depth=12, heads=6, features=64, out_channels=[48, 96, 192, 384], taps=[2, 5, 8, 11]),
    "vitb": dict(dim=768, depth=12, heads=12, features=128, out_channels=[96, 192, 384, 768], taps=[2, 5, 8, 11]),
    "vitl": dict(dim=1024, depth=24, heads=16, features=256, out_channels=[256, 512, 1024, 1024], taps=[4, 11, 17, 23]),
}
PATCH = 14
INFER_LEN, OVERLAP, INTERP_LEN = 32, 10, 8          # video_depth.py:29-33
KEYFRAMES = [0, 12, 24, 25, 26, 27, 28, 29, 30, 31]


def infer_arch(sd):
    dim = sd["pretrained.cls_token"].shape[-1]
    return {384: "vits", 768: "vitb", 1024: "vitl"}[dim]


# ----------------------------------------------------------------------------- encoder (DINOv2)
def pos_embed_for(sd, gh, gw):
    """dinov2.py:179-210 interpolate_pos_encoding (offset 0.1, bicubic, no antialias)."""
    pe = sd["pretrained.pos_embed"].float()
    n = pe.shape[1] - 1
    side = int(math.sqrt(n))
    if gh * gw == n and gh == gw:
        return pe
    dim = pe.shape[-1]
    # NB the reference names the image dims (w, h) = x.shape[2:], i.e. "w" is the height.
    w0, h0 = gh + 0.1, gw + 0.1
    grid = pe[:, 1:].reshape(1, side, side, dim).permute(0, 3, 1, 2)
    grid = F.interpolate(grid, scale_factor=(w0 / side, h0 / side), mode="bicubic", antialias=False)
    assert grid.shape[-2] == gh and grid.shape[-1] == gw
    grid = grid.permute(0, 2, 3, 1).reshape(1, -1, dim)
    return torch.cat([pe[:, :1], grid], dim=1)


def vit_block(sd, pre, x, heads):
    """dinov2_layers/block.py:82-107 eval branch with attention.py:49-62 and mlp.py:35-41."""
    B, N, C = x.shape
    h = F.layer_norm(x, (C,), sd[pre + "norm1.weight"], sd[pre + "norm1.bias"], 1e-6)
    qkv = F.linear(h, sd[pre + "attn.qkv.weight"], sd[pre + "attn.qkv.bias"])
    qkv = qkv.reshape(B, N, 3, heads, C // heads).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0] * (C // heads) ** -0.5, qkv[1], qkv[2]
    a = (q @ k.transpose(-2, -1)).softmax(dim=-1)
    h = (a @ v).transpose(1, 2).reshape(B, N, C)
    h = F.linear(h, sd[pre + "attn.proj.weight"], sd[pre + "attn.proj.bias"])
    x = x + h * sd[pre + "ls1.gamma"]
    h = F.layer_norm(x, (C,), sd[pre + "norm2.weight"], sd[pre + "norm2.bias"], 1e-6)
    h = F.gelu(F.linear(h, sd[pre + "mlp.fc1.weight"], sd[pre + "mlp.fc1.bias"]))
    h = F.linear(h, sd[pre + "mlp.fc2.weight"], sd[pre + "mlp.fc2.bias"])
    return x + h * sd[pre + "ls2.gamma"]


def encoder_taps(sd, x, arch):
    """get_intermediate_layers(..., norm=True) -> 4 x [frames, gh*gw, D] patch tokens (dinov2.py:297-321)."""
    cfg = ARCH[arch]
    frames, _, H, W = x.shape
    gh, gw = H // PATCH, W // PATCH
    t = F.conv2d(x, sd["pretrained.patch_embed.proj.weight"], sd["pretrained.patch_embed.proj.bias"], stride=PATCH)
    t = t.flatten(2).transpose(1, 2)
    t = torch.cat([sd["pretrained.cls_token"].expand(frames, -1, -1), t], dim=1) + pos_embed_for(sd, gh, gw)
    outs = []
    for i in range(cfg["depth"]):
        t = vit_block(sd, f"pretrained.blocks.{i}.", t, cfg["heads"])
        if i in cfg["taps"]:
            n = F.layer_norm(t, (t.shape[-1],), sd["pretrained.norm.weight"], sd["pretrained.norm.bias"], 1e-6)
            outs.append(n[:, 1:])
    return outs


# ----------------------------------------------------------------------------- motion module
def temporal_attention(sd, pre, h, T, heads, cached=None):
    """motion_module.py:242-321. h: [(b f), d, c] normalised hidden states.  Returns (out, cache)."""
    bf, d, c = h.shape
    if cached is None:
        seq = h.reshape(bf // T, T, d, c).permute(0, 2, 1, 3).reshape(-1, T, c)   # (b d) f c
        cache_out, d_in = seq, 0
    else:
        seq = h.reshape(bf, 1, d, c).permute(0, 2, 1, 3).reshape(-1, 1, c)
        cache_out, d_in = seq, cached.shape[1]
        seq = torch.cat([cached, seq], dim=1)
    seq = seq + sd[pre + "pos_encoder.pe"][:, : seq.shape[1]]
    q = F.linear(seq[:, d_in:], sd[pre + "to_q.weight"])
    k = F.linear(seq, sd[pre + "to_k.weight"])
    v = F.linear(seq, sd[pre + "to_v.weight"])

    def split(t):
        return t.reshape(t.shape[0], t.shape[1], heads, c // heads).permute(0, 2, 1, 3)

    q, k, v = split(q), split(k), split(v)
    a = (q @ k.transpose(-1, -2) * (c // heads) ** -0.5).softmax(dim=-1)
    o = (a @ v).permute(0, 2, 1, 3).reshape(q.shape[0], -1, c)
    o = F.linear(o, sd[pre + "to_out.0.weight"], sd[pre + "to_out.0.bias"])
    f = o.shape[1]
    o = o.reshape(-1, d, f, c).permute(0, 2, 1, 3).reshape(-1, d, c)             # (b f) d c
    return o, cache_out


def motion_module(sd, idx, x, T, cached=None):
    """motion_module.py:102-134 + :172-189. x: [(b f), C, h, w] -> same shape, list of 2 caches."""
    pre = f"head.motion_modules.{idx}.temporal_transformer."
    bf, C, hh, ww = x.shape
    h = F.group_norm(x, 32, sd[pre + "norm.weight"], sd[pre + "norm.bias"], 1e-6)
    h = h.permute(0, 2, 3, 1).reshape(bf, hh * ww, C)
    h = F.linear(h, sd[pre + "proj_in.weight"], sd[pre + "proj_in.bias"])
    blk = pre + "transformer_blocks.0."
    caches = []
    for a in range(2):
        n = F.layer_norm(h, (C,), sd[blk + f"norms.{a}.weight"], sd[blk + f"norms.{a}.bias"], 1e-5)
        o, cache = temporal_attention(sd, blk + f"attention_blocks.{a}.", n, T, 8,
                                      None if cached is None else cached[a])
        h = o + h
        caches.append(cache)
    n = F.layer_norm(h, (C,), sd[blk + "ff_norm.weight"], sd[blk + "ff_norm.bias"], 1e-5)
    g = F.linear(n, sd[blk + "ff.net.0.proj.weight"], sd[blk + "ff.net.0.proj.bias"])
    val, gate = g.chunk(2, dim=-1)
    h = F.linear(val * F.gelu(gate), sd[blk + "ff.net.2.weight"], sd[blk + "ff.net.2.bias"]) + h
    h = F.linear(h, sd[pre + "proj_out.weight"], sd[pre + "proj_out.bias"])
    h = h.reshape(bf, hh, ww, C).permute(0, 3, 1, 2)
    return h + x, caches


# ----------------------------------------------------------------------------- DPT head
def _rcu(sd, pre, x):
    """util/blocks.py:68-91 (bn=False; the skip uses the un-rectified input)."""
    h = F.conv2d(F.relu(x), sd[pre + "conv1.weight"], sd[pre + "conv1.bias"], padding=1)
    h = F.conv2d(F.relu(h), sd[pre + "conv2.weight"], sd[pre + "conv2.bias"], padding=1)
    return h + x


def _fusion(sd, name, xs, size=None):
    """util/blocks.py:135-162."""
    pre = f"head.scratch.{name}."
    out = xs[0]
    if len(xs) == 2:
        out = out + _rcu(sd, pre + "resConfUnit1.", xs[1])
    out = _rcu(sd, pre + "resConfUnit2.", out)
    if size is None:
        out = F.interpolate(out, scale_factor=2, mode="bilinear", align_corners=True)
    else:
        out = F.interpolate(out, size=size, mode="bilinear", align_corners=True)
    return F.conv2d(out, sd[pre + "out_conv.weight"], sd[pre + "out_conv.bias"])


def head_forward(sd, taps, gh, gw, T, cached=None):
    """dpt_temporal.py:53-125.  taps: 4 x [frames, gh*gw, D].  cached: list of 8 tensors or None.
    Returns (depth [frames,1,14gh,14gw], list of 8 caches)."""
    layers = []
    for i, t in enumerate(taps):
        x = t.permute(0, 2, 1).reshape(t.shape[0], t.shape[-1], gh, gw)
        x = F.conv2d(x, sd[f"head.projects.{i}.weight"], sd[f"head.projects.{i}.bias"])
        if i == 0:
            x = F.conv_transpose2d(x, sd["head.resize_layers.0.weight"], sd["head.resize_layers.0.bias"], stride=4)
        elif i == 1:
            x = F.conv_transpose2d(x, sd["head.resize_layers.1.weight"], sd["head.resize_layers.1.bias"], stride=2)
        elif i == 3:
            x = F.conv2d(x, sd["head.resize_layers.3.weight"], sd["head.resize_layers.3.bias"], stride=2, padding=1)
        layers.append(x)
    l1, l2, l3, l4 = layers
    c = (lambda j: None) if cached is None else (lambda j: cached[2 * j: 2 * j + 2])
    l3, h0 = motion_module(sd, 0, l3, T, c(0))
    l4, h1 = motion_module(sd, 1, l4, T, c(1))
    r1 = F.conv2d(l1, sd["head.scratch.layer1_rn.weight"], padding=1)
    r2 = F.conv2d(l2, sd["head.scratch.layer2_rn.weight"], padding=1)
    r3 = F.conv2d(l3, sd["head.scratch.layer3_rn.weight"], padding=1)
    r4 = F.conv2d(l4, sd["head.scratch.layer4_rn.weight"], padding=1)
    p4 = _fusion(sd, "refinenet4", [r4], size=r3.shape[2:])
    p4, h2 = motion_module(sd, 2, p4, T, c(2))
    p3 = _fusion(sd, "refinenet3", [p4, r3], size=r2.shape[2:])
    p3, h3 = motion_module(sd, 3, p3, T, c(3))
    p2 = _fusion(sd, "refinenet2", [p3, r2], size=r1.shape[2:])
    p1 = _fusion(sd, "refinenet1", [p2, r1])
    o = F.conv2d(p1, sd["head.scratch.output_conv1.weight"], sd["head.scratch.output_conv1.bias"], padding=1)
    o = F.interpolate(o, (gh * PATCH, gw * PATCH), mode="bilinear", align_corners=True)
    o = F.relu(F.conv2d(o, sd["head.scratch.output_conv2.0.weight"], sd["head.scratch.output_conv2.0.bias"], padding=1))
    o = F.relu(F.conv2d(o, sd["head.scratch.output_conv2.2.weight"], sd["head.scratch.output_conv2.2.bias"]))
    return o, h0 + h1 + h2 + h3


def forward(sd, x, arch=None, cached=None, return_cache=False):
    """VideoDepthAnything.forward (video_depth.py:61-68): x [B,T,3,H,W] fp32 -> [B,T,H,W]."""
    arch = arch or infer_arch(sd)
    B, T, _, H, W = x.shape
    assert B == 1, "the drivers only ever use B=1"
    taps = encoder_taps(sd, x.flatten(0, 1), arch)
    depth, caches = head_forward(sd, taps, H // PATCH, W // PATCH, T, cached)
    depth = F.relu(F.interpolate(depth, size=(H, W), mode="bilinear", align_corners=True))
    depth = depth.squeeze(1).unflatten(0, (B, T))
    return (depth, caches) if return_cache else depth


# ----------------------------------------------------------------------------- drivers
def net_input_size(h0, w0, input_size=518):
    """video_depth.py:71-75 aspect guard + util/transform.py:51-107 (lower_bound, keep aspect, x14)."""
    ratio = max(h0, w0) / min(h0, w0)
    if ratio > 1.78:
        input_size = int(input_size * 1.777 / ratio)
        input_size = round(input_size / 14) * 14
    sh, sw = input_size / h0, input_size / w0
    s = sw if sw > sh else sh                      # lower_bound keeps the larger scale

    def mult(x):
        y = int(np.round(x / 14) * 14)
        if y < input_size:                         # min_val = target size (transform.py:95-97)
            y = int(np.ceil(x / 14) * 14)
        return y

    return mult(s * h0), mult(s * w0)


def preprocess_frame(frame_u8, nh, nw):
    """util/transform.py:109-157: /255, cv2 INTER_CUBIC resize, (x-mean)/std in float64 -> f32, CHW."""
    import cv2

    img = frame_u8.astype(np.float32) / 255.0
    img = cv2.resize(img, (nw, nh), interpolation=cv2.INTER_CUBIC)
    img = (img - np.array([0.485, 0.456, 0.406])) / np.array([0.229, 0.224, 0.225])
    return np.ascontiguousarray(np.transpose(img, (2, 0, 1)).astype(np.float32))


def window_frame_indices(n_frames):
    """Closed form of the window composition (video_depth.py:92-115)."""
    step = INFER_LEN - OVERLAP
    clip = lambda i: min(i, n_frames - 1)
    wins = []
    for k, start in enumerate(range(0, n_frames, step)):
        cur = [clip(start + i) for i in range(INFER_LEN)]
        if k > 0:
            prev = wins[-1]
            cur[:OVERLAP] = [prev[j] for j in KEYFRAMES]
        wins.append(cur)
    return wins


def scale_and_shift(pred, target):
    """utils/util.py:40-62 with mask == 1: float32 sums, 2x2 normal equations by determinant."""
    pred = pred.astype(np.float32)
    target = target.astype(np.float32)
    mask = np.ones_like(pred, dtype=np.float32)
    a00 = np.sum(mask * pred * pred)
    a01 = np.sum(mask * pred)
    a11 = np.sum(mask)
    b0 = np.sum(mask * pred * target)
    b1 = np.sum(mask * target)
    det = a00 * a11 - a01 * a01
    if det != 0:
        return (a11 * b0 - a01 * b1) / det, (-a01 * b0 + a00 * b1) / det
    return 1, 0


def scale_and_shift_masked(prediction, target, mask, scale_only=False):
    """utils/util.py:16-62 as written (float32 sums over mask * ...): `compute_scale_and_shift`."""
    prediction, target, mask = (a.astype(np.float32) for a in (prediction, target, mask))
    a00 = np.sum(mask * prediction * prediction)
    a01 = np.sum(mask * prediction)
    a11 = np.sum(mask)
    b0 = np.sum(mask * prediction * target)
    if scale_only:                      # util.py:23-38
        return b0 / (a00 + 1e-6), 0
    b1 = np.sum(mask * target)
    det = a00 * a11 - a01 * a01
    if det != 0:
        return (a11 * b0 - a01 * b1) / det, (-a01 * b0 + a00 * b1) / det
    return 1, 0


def interpolate_frames(pre, post):
    """utils/util.py:65-74: linear cross-fade of two equally long frame lists."""
    n = len(pre)
    step = (1.0 - 0.0) / (n - 1)
    w = [0.0] + [i * step for i in range(1, n - 1)] + [1.0]
    return [pre[i] * (1 - w[i]) + post[i] * w[i] for i in range(n)]


def depth_to_uint8(frames):
    """utils/dc_utils.py:75-79: clip-global min/max, float32 arithmetic, truncating cast."""
    d_min, d_max = frames.min(), frames.max()
    return ((frames - d_min) / (d_max - d_min) * 255).astype(np.uint8)


def normalize_depth(depth, near_value, far_value, invert_depth):
    """external_engine/engine.py:1099-1141 for one frame (the engine calls it per frame, :402 / :785)."""
    lo, hi = depth.min(), depth.max()
    if hi > lo:
        out = (depth - lo) / (hi - lo)
        if far_value > near_value:
            out = near_value + (far_value - near_value) * out
        else:
            out = far_value + (near_value - far_value) * out
    else:
        out = np.full_like(depth, near_value)
    if invert_depth:
        out = far_value - (out - near_value)
    return out


def align_windows(depth_list, metric=False):
    """video_depth.py:120-160: sequential scale/shift alignment + 8-frame cross-fade."""
    aligned, ref = [], []
    n_align = OVERLAP - INTERP_LEN
    kf = KEYFRAMES[:n_align]
    for base in range(0, len(depth_list), INFER_LEN):
        if not aligned:
            aligned += depth_list[:INFER_LEN]
            ref = [depth_list[base + j] for j in kf]
            continue
        cur = [depth_list[base + i] for i in range(n_align)]
        s, t = (1.0, 0.0) if metric else scale_and_shift(np.concatenate(cur), np.concatenate(ref))
        fix = lambda d: _affine_clamp(d, s, t)
        pre = aligned[-INTERP_LEN:]
        post = [fix(depth_list[base + i]) for i in range(n_align, OVERLAP)]
        w = [0.0] + [i * (1.0 / (INTERP_LEN - 1)) for i in range(1, INTERP_LEN - 1)] + [1.0]   # util.py:65-74
        aligned[-INTERP_LEN:] = [pre[i] * (1 - w[i]) + post[i] * w[i] for i in range(INTERP_LEN)]
        aligned += [fix(depth_list[base + i]) for i in range(OVERLAP, INFER_LEN)]
        ref = ref[:1] + [fix(depth_list[base + j]) for j in kf[1:]]
    return aligned


def _affine_clamp(d, s, t):
    out = d * s + t
    out[out < 0] = 0
    return out


@torch.no_grad()
def infer_video_depth(sd, frames, input_size=518, metric=False, arch=None, device="cpu"):
    """video_depth.py:70-162 with fp32=True semantics. frames uint8 [N,H0,W0,3] -> float32 [N,H0,W0]."""
    n, h0, w0 = frames.shape[:3]
    nh, nw = net_input_size(h0, w0, input_size)
    depth_list = []
    for idx in window_frame_indices(n):
        x = torch.from_numpy(np.stack([preprocess_frame(frames[i], nh, nw) for i in idx]))[None].to(device)
        d = forward(sd, x, arch)
        d = F.interpolate(d.flatten(0, 1).unsqueeze(1).float(), size=(h0, w0), mode="bilinear", align_corners=True)
        depth_list += [d[i, 0].cpu().numpy() for i in range(d.shape[0])]
    return np.stack(align_windows(depth_list, metric)[:n], axis=0)


def upscale_to_original(depth, original_height, original_width):
    """external_engine/video_depth_anything/video_depth.py:181-203: per-frame bilinear (align_corners=True)
    upscale of the aligned maps [N,H0,W0] to the caller's original plate size."""
    t = torch.from_numpy(np.ascontiguousarray(depth)).unsqueeze(1).float()
    return F.interpolate(t, size=(original_height, original_width), mode="bilinear", align_corners=True)[:, 0].numpy()


class StreamOracle:
    """video_depth_stream.py:77-161 (fp32): one frame in, one depth map out, 8-tensor cache per frame."""

    def __init__(self, sd, arch=None, device="cpu"):
        self.sd, self.arch, self.device = sd, arch or infer_arch(sd), device
        self.caches, self.id, self.size = [], -1, None

    @torch.no_grad()
    def step(self, frame, input_size=518):
        self.id += 1
        h0, w0 = frame.shape[:2]
        if self.size is None:
            self.size = (h0, w0, *net_input_size(h0, w0, input_size))
        assert (h0, w0) == self.size[:2]
        nh, nw = self.size[2:]
        x = torch.from_numpy(preprocess_frame(frame, nh, nw))[None, None].to(self.device)
        if not self.caches:
            d, cache = forward(self.sd, x, self.arch, return_cache=True)
            self.caches = [cache] * INFER_LEN
        else:
            cur = self.caches[0:2] + self.caches[-INFER_LEN + 3:]
            assert len(cur) == INFER_LEN - 1
            cat = [torch.cat([c[i] for c in cur], dim=1) for i in range(8)]
            d, cache = forward(self.sd, x, self.arch, cached=cat, return_cache=True)
            self.caches.append(cache)
        if self.id + INFER_LEN > 41 + 1:           # gap == 41, video_depth_stream.py:58-59,155-159
            del self.caches[1]
        d = F.interpolate(d.flatten(0, 1).unsqueeze(1).float(), size=(h0, w0), mode="bilinear", align_corners=True)
        return d[-1, 0].cpu().numpy()

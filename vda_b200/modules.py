"""Parameter containers with exactly the reference's state_dict schema (SURVEY.md 8b "Weights").

These nn.Modules only OWN parameters (so load_state_dict(strict=True) of a reference checkpoint,
.to(device), .eval() behave as in the reference); none of them implements forward -- all arithmetic
goes through vda_b200.engine, which reads the parameters once and repacks them for the kernels.
Initialisation follows the reference so a fresh module is statistically the same model:
dinov2.py:126-132,332-336 (trunc-normal 0.02 Linear weights, zero bias, pos_embed 0.02, cls 1e-6,
LayerScale 1.0), torch defaults elsewhere, zeroed motion-module proj_out (motion_module.py:57-58).
"""
import math

import torch
from torch import nn

ARCH = {
    "vits": dict(dim=384, depth=12, heads=6, taps=(2, 5, 8, 11)),
    "vitb": dict(dim=768, depth=12, heads=12, taps=(2, 5, 8, 11)),
    "vitl": dict(dim=1024, depth=24, heads=16, taps=(4, 11, 17, 23)),
}
PATCH = 14
POS_GRID = 37  # 518 / 14


class _Holder(nn.Module):
    def forward(self, *a, **k):  # pragma: no cover - guards against accidental eager use
        raise RuntimeError("vda_b200 modules are parameter containers; run them through vda_b200.engine")


class _Gamma(_Holder):
    def __init__(self, dim, init):
        super().__init__()
        self.gamma = nn.Parameter(init * torch.ones(dim))


class _Attn(_Holder):
    def __init__(self, dim):
        super().__init__()
        self.qkv = nn.Linear(dim, 3 * dim)
        self.proj = nn.Linear(dim, dim)


class _Mlp(_Holder):
    def __init__(self, dim):
        super().__init__()
        self.fc1 = nn.Linear(dim, 4 * dim)
        self.fc2 = nn.Linear(4 * dim, dim)


class _Block(_Holder):
    def __init__(self, dim):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, eps=1e-6)
        self.attn = _Attn(dim)
        self.ls1 = _Gamma(dim, 1.0)
        self.norm2 = nn.LayerNorm(dim, eps=1e-6)
        self.mlp = _Mlp(dim)
        self.ls2 = _Gamma(dim, 1.0)


class _PatchEmbed(_Holder):
    def __init__(self, dim):
        super().__init__()
        self.proj = nn.Conv2d(3, dim, kernel_size=PATCH, stride=PATCH)


class DinoEncoder(_Holder):
    """`pretrained.*` subtree (dinov2.py:44-132 with DINOv2(model_name) settings, :398-415)."""

    def __init__(self, arch):
        super().__init__()
        cfg = ARCH[arch]
        dim = cfg["dim"]
        self.embed_dim = dim
        self.num_heads = cfg["heads"]
        self.patch_embed = _PatchEmbed(dim)
        self.cls_token = nn.Parameter(torch.zeros(1, 1, dim))
        self.pos_embed = nn.Parameter(torch.zeros(1, POS_GRID * POS_GRID + 1, dim))
        self.blocks = nn.ModuleList([_Block(dim) for _ in range(cfg["depth"])])
        self.norm = nn.LayerNorm(dim, eps=1e-6)
        self.mask_token = nn.Parameter(torch.zeros(1, dim))
        nn.init.trunc_normal_(self.pos_embed, std=0.02)
        nn.init.normal_(self.cls_token, std=1e-6)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                nn.init.zeros_(m.bias)


class _PE(_Holder):
    def __init__(self, dim, max_len):
        super().__init__()
        pos = torch.arange(max_len).unsqueeze(1)
        div = torch.exp(torch.arange(0, dim, 2) * (-math.log(10000.0) / dim))
        pe = torch.zeros(1, max_len, dim)
        pe[0, :, 0::2] = torch.sin(pos * div)
        pe[0, :, 1::2] = torch.cos(pos * div)
        self.register_buffer("pe", pe)


class _TemporalAttn(_Holder):
    def __init__(self, dim, max_len):
        super().__init__()
        self.to_q = nn.Linear(dim, dim, bias=False)
        self.to_k = nn.Linear(dim, dim, bias=False)
        self.to_v = nn.Linear(dim, dim, bias=False)
        self.to_out = nn.ModuleList([nn.Linear(dim, dim), nn.Dropout(0.0)])
        self.pos_encoder = _PE(dim, max_len)


class _GEGLU(_Holder):
    def __init__(self, dim):
        super().__init__()
        self.proj = nn.Linear(dim, 8 * dim)


class _FF(_Holder):
    def __init__(self, dim):
        super().__init__()
        self.net = nn.ModuleList([_GEGLU(dim), nn.Dropout(0.0), nn.Linear(4 * dim, dim)])


class _TemporalBlock(_Holder):
    def __init__(self, dim, max_len):
        super().__init__()
        self.attention_blocks = nn.ModuleList([_TemporalAttn(dim, max_len) for _ in range(2)])
        self.norms = nn.ModuleList([nn.LayerNorm(dim) for _ in range(2)])
        self.ff = _FF(dim)
        self.ff_norm = nn.LayerNorm(dim)


class _TemporalTransformer(_Holder):
    def __init__(self, dim, max_len):
        super().__init__()
        self.norm = nn.GroupNorm(32, dim, eps=1e-6, affine=True)
        self.proj_in = nn.Linear(dim, dim)
        self.transformer_blocks = nn.ModuleList([_TemporalBlock(dim, max_len)])
        self.proj_out = nn.Linear(dim, dim)
        nn.init.zeros_(self.proj_out.weight)
        nn.init.zeros_(self.proj_out.bias)


class _TemporalModule(_Holder):
    def __init__(self, dim, max_len):
        super().__init__()
        self.temporal_transformer = _TemporalTransformer(dim, max_len)


class _RCU(_Holder):
    def __init__(self, f):
        super().__init__()
        self.conv1 = nn.Conv2d(f, f, 3, padding=1)
        self.conv2 = nn.Conv2d(f, f, 3, padding=1)


class _Fusion(_Holder):
    def __init__(self, f):
        super().__init__()
        self.out_conv = nn.Conv2d(f, f, 1)
        self.resConfUnit1 = _RCU(f)
        self.resConfUnit2 = _RCU(f)


class _Scratch(_Holder):
    def __init__(self, oc, f):
        super().__init__()
        for i in range(4):
            setattr(self, f"layer{i + 1}_rn", nn.Conv2d(oc[i], f, 3, padding=1, bias=False))
        for i in range(1, 5):
            setattr(self, f"refinenet{i}", _Fusion(f))
        self.output_conv1 = nn.Conv2d(f, f // 2, 3, padding=1)
        self.output_conv2 = nn.Sequential(nn.Conv2d(f // 2, 32, 3, padding=1), nn.ReLU(True), nn.Conv2d(32, 1, 1),
                                          nn.ReLU(True), nn.Identity())


class DPTTemporalHead(_Holder):
    """`head.*` subtree (dpt.py:47-124, dpt_temporal.py:23-51)."""

    def __init__(self, in_channels, features, out_channels, num_frames):
        super().__init__()
        oc = list(out_channels)
        self.features, self.out_channels = features, oc
        self.projects = nn.ModuleList([nn.Conv2d(in_channels, c, 1) for c in oc])
        self.resize_layers = nn.ModuleList([
            nn.ConvTranspose2d(oc[0], oc[0], 4, stride=4), nn.ConvTranspose2d(oc[1], oc[1], 2, stride=2),
            nn.Identity(), nn.Conv2d(oc[3], oc[3], 3, stride=2, padding=1)])
        self.scratch = _Scratch(oc, features)
        self.motion_modules = nn.ModuleList([_TemporalModule(c, num_frames) for c in (oc[2], oc[3], features, features)])

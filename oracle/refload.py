"""ORACLE support (test infrastructure): imports the real reference from /root/reference (build
container only -- the path does not exist on the GPU box) and builds seeded weights.

seeded_state_dict() is deterministic torch-CPU RNG, so the GPU box regenerates the very same
weights without needing the reference.  The perturbation makes every fused epilogue observable:
the reference zero-initialises each motion module's proj_out (motion_module.py:57-58) and starts
LayerScale at 1, biases at 0 (dinov2.py:332-336, :409), which would leave those paths unpinned.
"""
import os
import sys
import types

import torch

REF_ROOT = "/root/reference"


def reference_available():
    return os.path.isdir(os.path.join(REF_ROOT, "video_depth_anything"))


def import_reference():
    """Returns (offline VideoDepthAnything class, streaming class) of the unmodified reference."""
    if "easydict" not in sys.modules:
        mod = types.ModuleType("easydict")

        class EasyDict(dict):
            __getattr__ = dict.__getitem__
            __setattr__ = dict.__setitem__

        mod.EasyDict = EasyDict
        sys.modules["easydict"] = mod
    # the reference's own packages are called video_depth_anything / utils, like this repo's drop-in
    # shims: load them under the reference root with a clean module table, then restore ours.
    saved = {k: v for k, v in sys.modules.items() if k.split(".")[0] in ("video_depth_anything", "utils")}
    for k in saved:
        del sys.modules[k]
    # the reference's packages have no __init__.py (namespace packages), so any regular package of the
    # same name further down sys.path (this repo's shims) would win: hide those entries meanwhile.
    old_path = list(sys.path)
    sys.path[:] = [REF_ROOT] + [p for p in old_path
                                if not os.path.isfile(os.path.join(p or ".", "video_depth_anything", "__init__.py"))]
    try:
        import contextlib
        import io

        with contextlib.redirect_stdout(io.StringIO()):
            from video_depth_anything.video_depth import VideoDepthAnything as Offline
            from video_depth_anything.video_depth_stream import VideoDepthAnything as Stream
    finally:
        sys.path[:] = old_path
        for k in [k for k in sys.modules if k.split(".")[0] in ("video_depth_anything", "utils")]:
            sys.modules["_vda_reference_." + k] = sys.modules.pop(k)
        sys.modules.update(saved)
    return Offline, Stream


def import_reference_utils():
    """(compute_scale_and_shift, get_interpolate_frames, normalize_depth) of the unmodified reference.
    utils/util.py is loaded by file path (this repo has a `utils` package of its own); normalize_depth is a method
    of the job engine class (external_engine/engine.py:1099-1141) whose module needs a GUI / job stack to import,
    so the method's own source is compiled by itself and called with self = None."""
    import ast
    import importlib.util
    import logging

    import numpy as np

    spec = importlib.util.spec_from_file_location("_vda_reference_utils_util", os.path.join(REF_ROOT, "utils", "util.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    src = open(os.path.join(REF_ROOT, "external_engine", "engine.py")).read()
    fn = next(n for n in ast.walk(ast.parse(src)) if isinstance(n, ast.FunctionDef) and n.name == "normalize_depth")
    ns = {"np": np, "logger": logging.getLogger("reference"), "print": lambda *a, **k: None}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "engine.py:normalize_depth", "exec"), ns)
    return mod.compute_scale_and_shift, mod.get_interpolate_frames, lambda *a: ns["normalize_depth"](None, *a)


def state_dict_shapes(arch):
    """Key -> shape of the reference state_dict, derived from the architecture table (no reference needed)."""
    from .vda_oracle import ARCH

    c = ARCH[arch]
    D, F_, oc = c["dim"], c["features"], c["out_channels"]
    s = {"pretrained.cls_token": (1, 1, D), "pretrained.pos_embed": (1, 1370, D), "pretrained.mask_token": (1, D),
         "pretrained.patch_embed.proj.weight": (D, 3, 14, 14), "pretrained.patch_embed.proj.bias": (D,)}
    for i in range(c["depth"]):
        p = f"pretrained.blocks.{i}."
        s.update({p + "norm1.weight": (D,), p + "norm1.bias": (D,), p + "attn.qkv.weight": (3 * D, D),
                  p + "attn.qkv.bias": (3 * D,), p + "attn.proj.weight": (D, D), p + "attn.proj.bias": (D,),
                  p + "ls1.gamma": (D,), p + "norm2.weight": (D,), p + "norm2.bias": (D,),
                  p + "mlp.fc1.weight": (4 * D, D), p + "mlp.fc1.bias": (4 * D,),
                  p + "mlp.fc2.weight": (D, 4 * D), p + "mlp.fc2.bias": (D,), p + "ls2.gamma": (D,)})
    s.update({"pretrained.norm.weight": (D,), "pretrained.norm.bias": (D,)})
    for i in range(4):
        s[f"head.projects.{i}.weight"] = (oc[i], D, 1, 1)
        s[f"head.projects.{i}.bias"] = (oc[i],)
    s.update({"head.resize_layers.0.weight": (oc[0], oc[0], 4, 4), "head.resize_layers.0.bias": (oc[0],),
              "head.resize_layers.1.weight": (oc[1], oc[1], 2, 2), "head.resize_layers.1.bias": (oc[1],),
              "head.resize_layers.3.weight": (oc[3], oc[3], 3, 3), "head.resize_layers.3.bias": (oc[3],)})
    for i in range(4):
        s[f"head.scratch.layer{i + 1}_rn.weight"] = (F_, oc[i], 3, 3)
    for r in range(1, 5):
        p = f"head.scratch.refinenet{r}."
        s[p + "out_conv.weight"], s[p + "out_conv.bias"] = (F_, F_, 1, 1), (F_,)
        for u in (1, 2):
            for cv in (1, 2):
                s[p + f"resConfUnit{u}.conv{cv}.weight"] = (F_, F_, 3, 3)
                s[p + f"resConfUnit{u}.conv{cv}.bias"] = (F_,)
    s.update({"head.scratch.output_conv1.weight": (F_ // 2, F_, 3, 3), "head.scratch.output_conv1.bias": (F_ // 2,),
              "head.scratch.output_conv2.0.weight": (32, F_ // 2, 3, 3), "head.scratch.output_conv2.0.bias": (32,),
              "head.scratch.output_conv2.2.weight": (1, 32, 1, 1), "head.scratch.output_conv2.2.bias": (1,)})
    for m, C in enumerate([oc[2], oc[3], F_, F_]):
        p = f"head.motion_modules.{m}.temporal_transformer."
        s.update({p + "norm.weight": (C,), p + "norm.bias": (C,), p + "proj_in.weight": (C, C),
                  p + "proj_in.bias": (C,), p + "proj_out.weight": (C, C), p + "proj_out.bias": (C,)})
        b = p + "transformer_blocks.0."
        for a in range(2):
            q = b + f"attention_blocks.{a}."
            s.update({q + "to_q.weight": (C, C), q + "to_k.weight": (C, C), q + "to_v.weight": (C, C),
                      q + "to_out.0.weight": (C, C), q + "to_out.0.bias": (C,), q + "pos_encoder.pe": (1, 32, C),
                      b + f"norms.{a}.weight": (C,), b + f"norms.{a}.bias": (C,)})
        s.update({b + "ff.net.0.proj.weight": (8 * C, C), b + "ff.net.0.proj.bias": (8 * C,),
                  b + "ff.net.2.weight": (C, 4 * C), b + "ff.net.2.bias": (C,),
                  b + "ff_norm.weight": (C,), b + "ff_norm.bias": (C,)})
    return s


def sinusoid_pe(C, max_len=32):
    """motion_module.py:192-210."""
    import math

    pos = torch.arange(max_len).unsqueeze(1)
    div = torch.exp(torch.arange(0, C, 2) * (-math.log(10000.0) / C))
    pe = torch.zeros(1, max_len, C)
    pe[0, :, 0::2] = torch.sin(pos * div)
    pe[0, :, 1::2] = torch.cos(pos * div)
    return pe


def seeded_state_dict(arch, seed=0):
    """Deterministic random weights with reference key names and every fused term non-trivial."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for k, shape in state_dict_shapes(arch).items():
        leaf = k.rsplit(".", 1)[-1]
        if leaf == "pe":
            sd[k] = sinusoid_pe(shape[-1])
        elif leaf == "gamma":
            sd[k] = 1.0 + 0.2 * torch.randn(shape, generator=g)
        elif leaf == "weight" and len(shape) == 1:                       # LayerNorm / GroupNorm scale
            sd[k] = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif leaf == "bias":
            sd[k] = 0.02 * torch.randn(shape, generator=g)
        elif k.endswith("cls_token") or k.endswith("pos_embed") or k.endswith("mask_token"):
            sd[k] = 0.02 * torch.randn(shape, generator=g)
        else:
            fan_in = 1
            for d in shape[1:]:
                fan_in *= d
            if "resize_layers.0" in k or "resize_layers.1" in k:          # ConvTranspose: [Cin, Cout, k, k]
                fan_in = shape[0]
            sd[k] = torch.randn(shape, generator=g) * (fan_in ** -0.5)
    # keep the regressed depth strictly positive (abs-rel parity metric divides by it): the last 1x1
    # conv sums ReLU outputs, so non-negative weights + a positive bias cannot hit the final ReLU.
    sd["head.scratch.output_conv2.2.weight"] = sd["head.scratch.output_conv2.2.weight"].abs()
    sd["head.scratch.output_conv2.2.bias"] = torch.tensor([0.1])
    return sd

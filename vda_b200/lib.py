"""ctypes binding of libvda_b200.so (C ABI declared in include/vda_b200.h).

There is deliberately no fallback: if the library is missing or the device is not sm_100 the
import/first call raises.  torch is used only to own device memory and streams.
"""
import ctypes as C
import os

import torch

from . import build as _build

ACT_NONE, ACT_GELU, ACT_RELU, ACT_GEGLU = 0, 1, 2, 3
ST_PLAIN, ST_CONVT, ST_QKV, ST_HEADDOT = 0, 1, 2, 3


class Epilogue(C.Structure):
    _fields_ = [
        ("bias", C.c_void_p), ("scale", C.c_void_p), ("res_f32", C.c_void_p),
        ("res1", C.c_void_p), ("res2", C.c_void_p),
        ("out", C.c_void_p), ("out_relu", C.c_void_p), ("out_f32", C.c_void_p),
        ("ldo", C.c_int), ("act", C.c_int), ("store_mode", C.c_int), ("block_n", C.c_int),
        ("ct_s", C.c_int), ("ct_cout", C.c_int), ("ct_h", C.c_int), ("ct_w", C.c_int),
        ("vt", C.c_void_p),
        ("qkv_d", C.c_int), ("qkv_heads", C.c_int), ("qkv_s", C.c_int), ("qkv_spad", C.c_int),
        ("w2", C.c_void_p), ("b2", C.c_float), ("cta_pair", C.c_int), ("split", C.c_int), ("ld32", C.c_int),
    ]


class VdaError(RuntimeError):
    pass


_lib = None
_inited = set()

# launch accounting for bench.py: LAUNCHES counts kernels of this library enqueued so far;
# when PROFILE is a list, every tensor-core launch appends (kind, flops, start_event, end_event).
LAUNCHES = 0
PROFILE = None
NOTES = None   # optional list filled alongside PROFILE with a shape string per launch (scripts/prof_launches.py)


def _launched(n=1):
    global LAUNCHES
    LAUNCHES += n


class _timed:
    """Brackets one launch with CUDA events on the current stream when profiling is on."""

    def __init__(self, kind, flops, note=""):
        self.kind, self.flops, self.note = kind, flops, note

    def __enter__(self):
        if PROFILE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()

    def __exit__(self, *a):
        if PROFILE is not None:
            self.e1.record()
            PROFILE.append((self.kind, self.flops, self.e0, self.e1))
            if NOTES is not None:
                NOTES.append(self.note)


def load():
    """Load the in-tree library (rebuilding it if it is absent or was built from other sources) and return the
    ctypes handle."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.build()     # no-op when the in-tree binary carries the digest of the current sources
    lib = C.CDLL(path)
    lib.vda_last_error.restype = C.c_char_p
    lib.vda_build_digest.restype = C.c_char_p
    if lib.vda_build_digest().decode() != _build._digest():
        raise VdaError(f"{path} was built from other sources than vda_b200/csrc (stale binary); rebuild with "
                       "`python -m vda_b200.build --force`")
    _declare(lib)
    _lib = lib
    return lib


def _declare(lib):
    vp, i, f = C.c_void_p, C.c_int, C.c_float
    ep = C.POINTER(Epilogue)
    sig = {
        "vda_init": [i],
        "vda_num_sms": [],
        "vda_set_pdl": [i],
        "vda_gemm_bf16": [vp, i, vp, i, i, i, i, ep, vp],
        "vda_conv3x3_nhwc": [vp, i, vp, i, i, i, i, i, ep, vp],
        "vda_flash_attn_fwd": [vp, vp, vp, i, i, i, i, vp, vp],
        "vda_layernorm": [vp, i, vp, vp, vp, vp, vp, i, i, f, i, i, i, i, i, vp],
        "vda_groupnorm_nhwc": [vp, i, vp, vp, vp, vp, i, i, i, f, i, vp],
        "vda_groupnorm_nhwc_static": [vp, i, vp, vp, vp, i, i, i, f, i, vp],
        "vda_patch_im2col": [vp, vp, i, i, i, i, i, vp],
        "vda_add_cls_pos": [vp, vp, vp, vp, i, i, i, vp],
        "vda_bilinear_nhwc": [vp, i, vp, i, i, i, i, i, i, i, i, i, vp],
        "vda_bilinear_f32": [vp, vp, i, i, i, i, i, i, vp],
        "vda_im2col_s2": [vp, vp, i, i, i, i, vp],
        "vda_temporal_attn": [vp, i, vp, i, vp, i, vp, i, i, i, i, i, i, vp],
        "vda_temporal_attn_stream": [vp, vp, vp, i, vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, i, vp],
        "vda_temporal_attn_stream_f32": [vp, vp, vp, i, vp, vp, vp, vp, vp, vp, vp, i, i, i, i, i, i, i, vp],
        "vda_preprocess_u8": [vp, i, i, i, vp, i, i, vp],
        "vda_minmax_f32": [vp, C.c_longlong, vp, vp],
        "vda_depth_colorize": [vp, C.c_longlong, vp, vp, vp, i, vp],
        "vda_depth_normalize": [vp, vp, C.c_longlong, i, C.c_double, C.c_double, i, vp, vp],
        "vda_depth_tail": [vp, i, vp, vp, vp, f, vp, i, i, i, i, i, i, vp],
        "vda_attn_f32": [vp, i, i, vp, i, i, i, i, i, vp],
        "vda_temporal_attn_f32": [vp, vp, vp, i, vp, i, i, i, i, i, i, i, vp],
        "vda_lsq_accum": [vp, vp, C.c_longlong, vp, vp],
        "vda_lsq_accum_masked": [vp, vp, vp, C.c_longlong, vp, vp],
        "vda_blend_frames": [vp, vp, vp, C.c_longlong, i, vp],
        "vda_lsq_solve": [vp, vp, i, vp],
        "vda_affine_clamp": [vp, vp, C.c_longlong, vp, vp],
        "vda_affine_clamp_blend": [vp, vp, C.c_longlong, i, vp, vp],
    }
    for name, args in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = i
    return sig


def set_pdl(on):
    """Programmatic dependent launch for the launches that follow (csrc/common.cuh); returns the previous setting."""
    return bool(load().vda_set_pdl(int(bool(on))))


def exported_symbols():
    """Names every include/vda_b200.h declaration must resolve to (used by the CPU-side test)."""
    return sorted(_declare(load()).keys()) + ["vda_last_error", "vda_build_digest"]


def check(rc):
    if rc != 0:
        raise VdaError(f"libvda_b200 error {rc}: {load().vda_last_error().decode()}")


def init(device=None):
    """Initialise the library for a CUDA device (idempotent). Raises if no sm_100 GPU."""
    if not torch.cuda.is_available():
        raise VdaError("libvda_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    dev = torch.cuda.current_device() if device is None else torch.device(device).index or 0
    if dev not in _inited:
        check(load().vda_init(dev))
        _inited.add(dev)
    return dev


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def make_epilogue(bias=None, scale=None, res_f32=None, res1=None, res2=None, out=None, out_relu=None,
                  out_f32=None, ldo=0, act=ACT_NONE, store_mode=ST_PLAIN, block_n=0, ct=None, qkv=None,
                  w2=None, b2=0.0, cta_pair=0, split=0, ld32=0):
    e = Epilogue()
    e.bias, e.scale, e.res_f32 = ptr(bias), ptr(scale), ptr(res_f32)
    e.res1, e.res2 = ptr(res1), ptr(res2)
    e.out, e.out_relu, e.out_f32 = ptr(out), ptr(out_relu), ptr(out_f32)
    e.ldo, e.act, e.store_mode, e.block_n = ldo, act, store_mode, block_n
    if ct is not None:
        e.ct_s, e.ct_cout, e.ct_h, e.ct_w = ct
    if qkv is not None:
        vt, d, heads, s, spad = qkv
        e.vt, e.qkv_d, e.qkv_heads, e.qkv_s, e.qkv_spad = ptr(vt), d, heads, s, spad
    e.w2, e.b2, e.cta_pair = ptr(w2), b2, cta_pair
    e.split, e.ld32 = split, ld32
    return e


def gemm(a, w, epi, M=None, N=None, K=None, lda=None, ldw=None):
    """epi(A[M,K] @ W[N,K]^T); a, w bf16 row-major 2-D CUDA tensors."""
    M = a.shape[0] if M is None else M
    K = a.shape[1] if K is None else K
    N = w.shape[0] if N is None else N
    lda = a.stride(0) if lda is None else lda
    ldw = w.stride(0) if ldw is None else ldw
    with _timed("gemm", 2.0 * M * N * K, f"gemm M={M} N={N} K={K} act={epi.act} st={epi.store_mode} r32={int(bool(epi.res_f32))} "
                                          f"o={int(bool(epi.out))} of={int(bool(epi.out_f32))} r1={int(bool(epi.res1))}"):
        check(load().vda_gemm_bf16(ptr(a), lda, ptr(w), ldw, M, N, K, C.byref(epi), stream_ptr()))
    _launched()


def conv3x3(x, w, epi, n, H, W, Cin, Cout, ldx=None):
    ldx = Cin if ldx is None else ldx
    with _timed("conv3x3", 2.0 * n * H * W * Cout * 9 * Cin, f"conv n={n} {H}x{W} {Cin}->{Cout} act={epi.act} r1={int(bool(epi.res1))} "
                                                             f"r2={int(bool(epi.res2))} orelu={int(bool(epi.out_relu))}"):
        check(load().vda_conv3x3_nhwc(ptr(x), ldx, ptr(w), n, H, W, Cin, Cout, C.byref(epi), stream_ptr()))
    _launched()


def flash_attn(qk, vt, out, frames, S, heads, s_pad, flags=None):
    """Exact online softmax.  flags (int32 [frames*heads*ceil(S/128)] zeros) only matter to experiment builds of the
    library (-DVDA_ATTN_EXPERIMENTS with VDA_ATTN_OPT=1: optimistic softmax + fix-up pass)."""
    with _timed("attn", 4.0 * frames * heads * S * S * 64):
        check(load().vda_flash_attn_fwd(ptr(qk), ptr(vt), ptr(out), frames, S, heads, s_pad, ptr(flags), stream_ptr()))
    _launched(1)


def depth_tail_supported(h, w, H, W, C):
    """True if vda_depth_tail takes this shape (host restatement of its argument check)."""
    if C not in (64, 128) or h <= 1 or w <= 1 or H < h or W < w:
        return False
    sy, sx = (h - 1) / (H - 1), (w - 1) / (W - 1)
    return int(17 * sy + 1e-3) + 3 <= 12 and int(9 * sx + 1e-3) + 3 <= 8


def depth_tail(x, w3, b3, w1, b1, out, n, h, w, H, W, C, ldx=None):
    """Fused bilinear upsample + 3x3 conv (C -> 32) + ReLU + 1x1 (32 -> 1) + ReLU; x NHWC bf16, out f32 [n,H,W]."""
    nbytes = n * (h * w * C * 2 + H * W * 4)
    with _timed("conv3x3", 2.0 * n * H * W * 32 * 9 * C, f"depth_tail n={n} {h}x{w}->{H}x{W} C={C} ({nbytes >> 20} MB algorithmic)"):
        check(load().vda_depth_tail(ptr(x), C if ldx is None else ldx, ptr(w3), ptr(b3), ptr(w1), b1, ptr(out), n, h, w, H, W,
                                    C, stream_ptr()))
    _launched()


def depth_colorize(depth, lut_rgb=None, grey=False):
    """depth f32 [...] on the GPU -> uint8 [..., 3] (or [...] if grey) with the clip-global min/max."""
    mm = torch.empty(2, dtype=torch.float32, device=depth.device)
    check(load().vda_minmax_f32(ptr(depth), depth.numel(), ptr(mm), stream_ptr()))
    out = torch.empty(tuple(depth.shape) + (() if grey else (3,)), dtype=torch.uint8, device=depth.device)
    check(load().vda_depth_colorize(ptr(depth), depth.numel(), ptr(mm), ptr(lut_rgb), ptr(out), int(grey), stream_ptr()))
    _launched(4)
    return out


def preprocess_u8(frames_u8, out, nh, nw):
    """frames_u8 uint8 [n,H0,W0,3] on the GPU -> out f32 [n,3,nh,nw]."""
    n, h0, w0, _ = frames_u8.shape
    check(load().vda_preprocess_u8(ptr(frames_u8), n, h0, w0, ptr(out), nh, nw, stream_ptr()))
    _launched()


def attn_f32(qkv, out, split, frames, S, heads, in_split):
    """qkv: split bf16 [frames*S, 3*in_split] (hi | lo | hi); out split bf16."""
    with _timed("attn", 4.0 * frames * heads * S * S * 64):
        check(load().vda_attn_f32(ptr(qkv), qkv.shape[1], in_split, ptr(out), out.shape[1], split, frames, S, heads,
                                  stream_ptr()))
    _launched()


def temporal_attn_f32(qkv, C_, out, split, hw, heads, Tq, Tk):
    """qkv f32 [T*hw, 3*C_] (q | k | v); out split bf16 [Tq*hw, 3*split]."""
    check(load().vda_temporal_attn_f32(ptr(qkv), ptr(qkv[:, C_:]), ptr(qkv[:, 2 * C_:]), qkv.shape[1], ptr(out),
                                       out.shape[1], split, hw, heads, C_ // heads, Tq, Tk, stream_ptr()))
    _launched()


def layernorm(x, gamma, beta, out, eps, rows=None, C=None, tokens_per_frame=0, skip=0, pe=None, out_nope=None,
              pe_rows_per_t=0, pe_T=0, split=0):
    rows = x.shape[0] if rows is None else rows
    C = x.shape[-1] if C is None else C
    nbytes = rows * C * (x.element_size() + 2 * (3 if split else 1) * (2 if out_nope is not None else 1))
    with _timed("hbm:layernorm", nbytes, f"layernorm rows={rows} C={C}"):   # for hbm:* kinds `flops` holds algorithmic bytes
        check(load().vda_layernorm(ptr(x), int(x.dtype == torch.float32), ptr(gamma), ptr(beta), ptr(out), ptr(out_nope),
                                   ptr(pe), rows, C, eps, tokens_per_frame, skip, pe_rows_per_t, pe_T, split, stream_ptr()))
    _launched(1)


def groupnorm_nhwc(x, gamma, beta, out, stats_ws, n, hw, C, eps, ldx=None, split=0):
    parts = 3 if split else 1
    with _timed("hbm:groupnorm", n * hw * C * 2 * (2 * (2 if split else 1) + parts), f"groupnorm n={n} hw={hw} C={C}"):
        check(load().vda_groupnorm_nhwc(ptr(x), C if ldx is None else ldx, ptr(gamma), ptr(beta), ptr(out), ptr(stats_ws),
                                        n, hw, C, eps, split, stream_ptr()))
    _launched(3)


GROUPNORM_STATIC_MAX_N = 64    # VDA_GROUPNORM_STATIC_MAX_N of include/vda_b200.h


def groupnorm_nhwc_static(x, gamma, beta, out, n, hw, C, eps, ldx=None, split=0):
    """GroupNorm whose scratch is the library's file-scope device array (no allocation anywhere; n <= 64)."""
    check(load().vda_groupnorm_nhwc_static(ptr(x), C if ldx is None else ldx, ptr(gamma), ptr(beta), ptr(out),
                                           n, hw, C, eps, split, stream_ptr()))
    _launched(3)


def patch_im2col(x, a, frames, H, W, ldk, split=0):
    check(load().vda_patch_im2col(ptr(x), ptr(a), frames, H, W, ldk, split, stream_ptr()))
    _launched(1)


def add_cls_pos(patch, cls, pos, tokens, frames, hw, D):
    check(load().vda_add_cls_pos(ptr(patch), ptr(cls), ptr(pos), ptr(tokens), frames, hw, D, stream_ptr()))
    _launched(1)


def bilinear_nhwc(x, out, n, h, w, H, W, C, ldx=None, ldo=None, split_in=0, split_out=0):
    parts = 3 if split_out else 1
    with _timed("hbm:bilinear", n * C * 2 * (h * w * (2 if split_in else 1) + H * W * parts), f"bilinear n={n} {h}x{w}->{H}x{W} C={C}"):
        check(load().vda_bilinear_nhwc(ptr(x), C if ldx is None else ldx, ptr(out), C if ldo is None else ldo,
                                       n, h, w, H, W, C, split_in, split_out, stream_ptr()))
    _launched(1)


def bilinear_f32(x, out, n, h, w, H, W, relu=False):
    check(load().vda_bilinear_f32(ptr(x), ptr(out), n, h, w, H, W, int(relu), stream_ptr()))
    _launched(1)


def im2col_s2(x, a, n, h, w, C):
    check(load().vda_im2col_s2(ptr(x), ptr(a), n, h, w, C, stream_ptr()))
    _launched(1)


def temporal_attn(q, ldq, k, ldk, v, ldv, out, ldo, hw, heads, head_dim, Tq, Tk):
    C_ = heads * head_dim
    with _timed("hbm:temporal_attn", hw * C_ * 2 * (2 * Tq + 2 * Tk), f"temporal_attn hw={hw} C={C_} T={Tk}"):
        check(load().vda_temporal_attn(ptr(q), ldq, ptr(k), ldk, ptr(v), ldv, ptr(out), ldo, hw, heads, head_dim, Tq, Tk,
                                       stream_ptr()))
    _launched(1)


def temporal_attn_stream(q, knew, vnew, ldn, kcache, vcache, slots, pe_q, pe_k, pe_v, out, ldo, hw, heads, head_dim, Tk):
    """kcache / vcache: [hw, cap, C] (the cap slots of a token are adjacent)."""
    assert kcache.dim() == 3 and kcache.shape[0] == hw and vcache.shape == kcache.shape
    check(load().vda_temporal_attn_stream(ptr(q), ptr(knew), ptr(vnew), ldn, ptr(kcache), ptr(vcache), ptr(slots),
                                          ptr(pe_q), ptr(pe_k), ptr(pe_v), ptr(out), ldo, hw, heads, head_dim, Tk,
                                          kcache.shape[1], stream_ptr()))
    _launched(1)


def temporal_attn_stream_f32(q, knew, vnew, ldn, kcache, vcache, slots, pe_q, pe_k, pe_v, out, split, hw, heads, head_dim,
                             Tk):
    """FP32-accuracy mode: q/knew/vnew/caches f32, out split bf16 [hw, 3*split]."""
    assert kcache.dim() == 3 and kcache.shape[0] == hw and vcache.shape == kcache.shape
    check(load().vda_temporal_attn_stream_f32(ptr(q), ptr(knew), ptr(vnew), ldn, ptr(kcache), ptr(vcache), ptr(slots),
                                              ptr(pe_q), ptr(pe_k), ptr(pe_v), ptr(out), out.shape[1], split, hw, heads,
                                              head_dim, Tk, kcache.shape[1], stream_ptr()))
    _launched(1)


def lsq_accum(pred, target, n, sums5, mask=None):
    """sums5 (f64, caller-zeroed) += {sum m p p, sum m p, sum m, sum m p t, sum m t}; mask None = all ones."""
    check(load().vda_lsq_accum_masked(ptr(pred), ptr(target), ptr(mask), n, ptr(sums5), stream_ptr()))
    _launched(1)


def lsq_solve(sums5, scale_shift, metric=False, reset=False):
    """{scale, shift} from the five sums (metric: {1, 0}); reset=True zeroes sums5 for the next window."""
    check(load().vda_lsq_solve(ptr(sums5), ptr(scale_shift), int(bool(metric)) | (2 if reset else 0), stream_ptr()))
    _launched(1)


def blend_frames(pre, post, out, per_frame, frames):
    check(load().vda_blend_frames(ptr(pre), ptr(post), ptr(out), per_frame, frames, stream_ptr()))
    _launched(1)


def depth_normalize(depth, near_value, far_value, invert_depth=False):
    """depth f32 [frames, ...] on the GPU -> f32 same shape, each frame mapped to [near, far]
    (external_engine/engine.py:1099-1141)."""
    d = depth.contiguous()
    frames = d.shape[0]
    out = torch.empty_like(d)
    ws = torch.empty(2 * frames, dtype=torch.float32, device=d.device)
    check(load().vda_depth_normalize(ptr(d), ptr(out), d[0].numel(), frames, float(near_value), float(far_value),
                                     int(bool(invert_depth)), ptr(ws), stream_ptr()))
    _launched(4)
    return out


def affine_clamp(src, dst, n, scale_shift):
    check(load().vda_affine_clamp(ptr(src), ptr(dst), n, ptr(scale_shift), stream_ptr()))
    _launched(1)


def affine_clamp_blend(pre, post, per_frame, frames, scale_shift):
    check(load().vda_affine_clamp_blend(ptr(pre), ptr(post), per_frame, frames, ptr(scale_shift), stream_ptr()))
    _launched(1)

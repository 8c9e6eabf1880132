"""Window executor: VideoDepthAnything.forward (video_depth.py:61-68) as a fixed sequence of
libvda_b200 kernel launches on the current CUDA stream.

Data layout: activations are row-major [rows, channels] -- tokens
[frame*S + s, D] in the encoder, NHWC pixels [((frame*h)+y)*w + x, C] in the head -- bf16, with the
two residual streams (ViT tokens, motion-module hidden states) kept in fp32.  torch only owns the
buffers (one workspace per input shape, reused for every window); no torch op runs per window.
"""
import os

import torch

from . import lib as L
from . import packing as P
from .modules import ARCH, PATCH, POS_GRID


def _pad(n, m):
    return (n + m - 1) // m * m


class Workspace:
    """Named device buffers, allocated on first use and reused for every window of the same shape."""

    def __init__(self, device):
        self.device = device
        self.bufs = {}

    def get(self, name, shape, dtype=torch.bfloat16, zero=False):
        key = (name, tuple(shape), dtype)
        t = self.bufs.get(key)
        if t is None:
            t = (torch.zeros if zero else torch.empty)(tuple(shape), dtype=dtype, device=self.device)
            self.bufs[key] = t
        return t

    def nbytes(self):
        return sum(t.numel() * t.element_size() for t in self.bufs.values())


class _Motion:
    """Packed weights of one TemporalModule (motion_module.py:68-134)."""

    def __init__(self, mod, dev):
        tt = mod.temporal_transformer
        blk = tt.transformer_blocks[0]
        self.C = tt.proj_in.weight.shape[0]
        self.gn_w, self.gn_b = P.vec(tt.norm.weight, dev), P.vec(tt.norm.bias, dev)
        self.w_in, self.b_in = P.linear(tt.proj_in.weight, tt.proj_in.bias, device=dev)
        self.w_out, self.b_out = P.linear(tt.proj_out.weight, tt.proj_out.bias, device=dev)
        self.attn = []
        for a, n in zip(blk.attention_blocks, blk.norms):
            wqkv, _ = P.linear(torch.cat([a.to_q.weight, a.to_k.weight, a.to_v.weight], 0), device=dev)
            wo, bo = P.linear(a.to_out[0].weight, a.to_out[0].bias, device=dev)
            pe = a.pos_encoder.pe.detach().float().reshape(-1, self.C).contiguous().to(dev)
            # streaming mode: K(n + pe) = K n + K pe, so the projected PE rows are added at read time
            proj_pe = [(pe @ w_.detach().float().to(dev).T).contiguous() for w_ in (a.to_q.weight, a.to_k.weight, a.to_v.weight)]
            self.attn.append(dict(ln_w=P.vec(n.weight, dev), ln_b=P.vec(n.bias, dev), wqkv=wqkv, wo=wo, bo=bo, pe=pe,
                                  pe_q=proj_pe[0], pe_k=proj_pe[1], pe_v=proj_pe[2]))
        self.ffn_w, self.ffn_b = P.vec(blk.ff_norm.weight, dev), P.vec(blk.ff_norm.bias, dev)
        self.w_geglu, self.b_geglu = P.geglu(blk.ff.net[0].proj.weight, blk.ff.net[0].proj.bias, dev)
        self.w_ff2, self.b_ff2 = P.linear(blk.ff.net[2].weight, blk.ff.net[2].bias, device=dev)


class _Fusion:
    def __init__(self, mod, dev):
        self.rcu = []
        for u in (mod.resConfUnit1, mod.resConfUnit2):
            w1, b1 = P.conv3x3(u.conv1.weight, u.conv1.bias, device=dev)
            w2, b2 = P.conv3x3(u.conv2.weight, u.conv2.bias, device=dev)
            self.rcu.append((w1, b1, w2, b2))
        self.w_out, self.b_out = P.linear(mod.out_conv.weight, mod.out_conv.bias, device=dev)


class Packed:
    """All kernel-layout weights of one VideoDepthAnything module."""

    def __init__(self, model, dev, split=False):
        with P.split_mode(split):
            self._pack(model, dev)

    def _pack(self, model, dev):
        enc, head = model.pretrained, model.head
        self.arch = model.encoder
        cfg = ARCH[self.arch]
        self.D, self.heads, self.taps = cfg["dim"], cfg["heads"], cfg["taps"]
        D = self.D
        wp = torch.zeros(D, 592, dtype=torch.float32, device=dev)
        wp[:, :588] = enc.patch_embed.proj.weight.detach().reshape(D, 588).to(dev).float()
        self.w_patch, self.b_patch = P.to_kernel(wp), P.vec(enc.patch_embed.proj.bias, dev)
        self.cls = P.vec(enc.cls_token, dev)
        self.pos_embed = enc.pos_embed.detach().float().to(dev)
        self.pos_cache = {}
        self.blocks = []
        for b in enc.blocks:
            wqkv, bqkv = P.linear(b.attn.qkv.weight, b.attn.qkv.bias, device=dev)
            wpr, bpr = P.linear(b.attn.proj.weight, b.attn.proj.bias, device=dev)
            w1, b1 = P.linear(b.mlp.fc1.weight, b.mlp.fc1.bias, device=dev)
            w2, b2 = P.linear(b.mlp.fc2.weight, b.mlp.fc2.bias, device=dev)
            self.blocks.append(dict(
                n1w=P.vec(b.norm1.weight, dev), n1b=P.vec(b.norm1.bias, dev), wqkv=wqkv, bqkv=bqkv, wproj=wpr, bproj=bpr,
                g1=P.vec(b.ls1.gamma, dev), n2w=P.vec(b.norm2.weight, dev), n2b=P.vec(b.norm2.bias, dev),
                w1=w1, b1=b1, w2=w2, b2=b2, g2=P.vec(b.ls2.gamma, dev)))
        self.norm_w, self.norm_b = P.vec(enc.norm.weight, dev), P.vec(enc.norm.bias, dev)
        # ---- head
        self.F = head.features
        self.oc = list(head.out_channels)
        self.ocp = [_pad(c, 64) for c in self.oc]      # channel counts as 3x3-conv inputs
        self.proj = [P.linear(m.weight, m.bias, device=dev) for m in head.projects]
        r0, r1, r3 = head.resize_layers[0], head.resize_layers[1], head.resize_layers[3]
        self.ct0 = P.conv_transpose(r0.weight, r0.bias, device=dev)
        self.ct1 = P.conv_transpose(r1.weight, r1.bias, device=dev)
        self.w_rs3, self.b_rs3 = P.conv3x3(r3.weight, r3.bias, cin_pad=8, device=dev)   # explicit im2col: no cin pad
        sc = head.scratch
        self.rn = [P.conv3x3(getattr(sc, f"layer{i + 1}_rn").weight, None, device=dev)[0] for i in range(4)]
        self.fusion = {i: _Fusion(getattr(sc, f"refinenet{i}"), dev) for i in range(1, 5)}
        self.F2p = _pad(self.F // 2, 64)
        self.w_oc1, self.b_oc1 = P.conv3x3(sc.output_conv1.weight, sc.output_conv1.bias, cout_pad=64, device=dev)
        self.w_oc2, self.b_oc2 = P.conv3x3(sc.output_conv2[0].weight, sc.output_conv2[0].bias, device=dev)
        self.w_last = P.vec(sc.output_conv2[2].weight, dev)
        self.b_last = float(sc.output_conv2[2].bias.detach().float().item())
        self.motion = [_Motion(m, dev) for m in head.motion_modules]

    def pos_for(self, gh, gw):
        """Positional table [1+gh*gw, D] f32 for a gh x gw grid (dinov2.py:179-210).  The bicubic
        resample of the 37x37 table for non-square inputs happens once per shape and is cached."""
        key = (gh, gw)
        if key not in self.pos_cache:
            pe = self.pos_embed
            if gh == POS_GRID and gw == POS_GRID:
                tab = pe[0]
            else:
                D = pe.shape[-1]
                grid = pe[:, 1:].reshape(1, POS_GRID, POS_GRID, D).permute(0, 3, 1, 2)
                grid = torch.nn.functional.interpolate(
                    grid, scale_factor=((gh + 0.1) / POS_GRID, (gw + 0.1) / POS_GRID), mode="bicubic", antialias=False)
                assert grid.shape[-2:] == (gh, gw)
                tab = torch.cat([pe[0, :1], grid.permute(0, 2, 3, 1).reshape(-1, D)], 0)
            self.pos_cache[key] = tab.contiguous()
        return self.pos_cache[key]


class Engine:
    """Executes the window forward for one model on one device."""

    def __init__(self, model, device, fp32=False):
        """fp32=True builds the FP32-accuracy executor (video_depth.py:108, `fp32` of infer_video_depth):
        the same launch sequence on split activations [hi | lo | hi] (3 bf16 parts per value) against
        weights packed [Wh | Wh | Wl], i.e. 3-term bf16 products with fp32 accumulation on the same
        tensor-core kernels, fp32 SIMT attention, and fp32 everywhere else (csrc/fp32path.cu)."""
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise L.VdaError("vda_b200 runs on sm_100a GPUs only; there is no CPU path")
        L.init(self.device)
        self.fp32 = bool(fp32)
        self.W = 3 if self.fp32 else 1  # bf16 parts per activation value
        with torch.cuda.device(self.device):
            self.w = Packed(model, self.device, split=self.fp32)
        self.ws = {}
        self.graphs = {}
        self._capture_stream = None
        self.launches_per_run = {}
        self.use_graph = True
        self._warm = set()
        self.tail_fused = os.environ.get("VDA_TAIL_FUSED", "1") == "1"
        # reassemble branches on side streams: 2 = always (default), 1 = streamed frames only, 0 = off (A/B knob)
        self.head_streams = int(os.environ.get("VDA_HEAD_STREAMS", "2"))
        self._side = None

    def _ep(self, width, **kw):
        """Epilogue of a launch whose bf16 outputs / residuals are `width` channels wide (x3 parts in
        FP32-accuracy mode); fp32 operands of the same launch keep pitch kw['ld32'] if given."""
        if self.fp32:
            return L.make_epilogue(ldo=width * 3, split=width, **kw)
        kw.pop("ld32", None)
        return L.make_epilogue(ldo=width, **kw)

    def _ws(self, key):
        if key not in self.ws:
            self.ws[key] = Workspace(self.device)
        return self.ws[key]

    # ------------------------------------------------------------------ encoder
    def encoder(self, T, H, W, ws):
        """Patch rows of a [T,3,H,W] window (gathered by _patch) -> 4 taps, each bf16 [T*gh*gw, D] (final-norm
        applied, cls dropped)."""
        w = self.w
        gh, gw = H // PATCH, W // PATCH
        hw, D, heads = gh * gw, w.D, w.heads
        S, M = hw + 1, T * (hw + 1)
        spad = _pad(S, 8)
        sp, Wd = self.fp32, self.W
        a = ws.get("patch_a", (T * hw, 592 * Wd))
        patch = ws.get("patch_f32", (T * hw, D), torch.float32)
        L.gemm(a, w.w_patch, L.make_epilogue(bias=w.b_patch, out_f32=patch, ldo=D), K=None if sp else 588)
        tok = ws.get("tokens", (M, D), torch.float32)
        L.add_cls_pos(patch, w.cls, w.pos_for(gh, gw), tok, T, hw, D)
        xn = ws.get("xn", (M, D * Wd))
        ao = ws.get("attn_out", (M, D * Wd))
        h1 = ws.get("mlp_h", (M, 4 * D * Wd))
        if sp:
            qkvs = ws.get("qkv_split", (M, 9 * D))   # q | k | v as [hi | lo | hi], parts 3D apart
        else:
            qk = ws.get("qk", (M, 2 * D))
            vt = ws.get("vt", (T * heads, 64, spad), zero=True)
        taps = []
        for i, b in enumerate(w.blocks):
            L.layernorm(tok, b["n1w"], b["n1b"], xn, 1e-6, C=D, split=int(sp))
            if sp:
                L.gemm(xn, b["wqkv"], self._ep(3 * D, bias=b["bqkv"], out=qkvs))
                L.attn_f32(qkvs, ao, D, T, S, heads, in_split=3 * D)
            else:
                L.gemm(xn, b["wqkv"], L.make_epilogue(bias=b["bqkv"], out=qk, ldo=2 * D, store_mode=L.ST_QKV,
                                                      qkv=(vt, D, heads, S, spad)))
                L.flash_attn(qk, vt, ao, T, S, heads, spad)
            L.gemm(ao, b["wproj"], L.make_epilogue(bias=b["bproj"], scale=b["g1"], res_f32=tok, out_f32=tok, ldo=D))
            L.layernorm(tok, b["n2w"], b["n2b"], xn, 1e-6, C=D, split=int(sp))
            L.gemm(xn, b["w1"], self._ep(4 * D, bias=b["b1"], out=h1, act=L.ACT_GELU))
            L.gemm(h1, b["w2"], L.make_epilogue(bias=b["b2"], scale=b["g2"], res_f32=tok, out_f32=tok, ldo=D))
            if i in w.taps:
                t = ws.get(f"tap{len(taps)}", (T * hw, D * Wd))
                L.layernorm(tok, w.norm_w, w.norm_b, t, 1e-6, C=D, tokens_per_frame=S, skip=1, split=int(sp))
                taps.append(t)
        return taps

    # ------------------------------------------------------------------ motion module
    def motion(self, idx, x, T, hw, ws, out, out_relu=None, stream_ctx=None):
        """x, out: NHWC bf16 [T*hw, C].  motion_module.py:102-134 / :172-189 / :242-321."""
        m = self.w.motion[idx]
        C, rows = m.C, T * hw
        sp, Wd = self.fp32, self.W
        tag = f"mm{idx}_"
        gn = ws.get(tag + "gn", (rows, C * Wd))
        stats = ws.get(tag + "gnstats", (96 * T,), torch.float64)
        L.groupnorm_nhwc(x, m.gn_w, m.gn_b, gn, stats, T, hw, C, 1e-6, ldx=C * Wd, split=C if sp else 0)
        h = ws.get(tag + "h", (rows, C), torch.float32)
        L.gemm(gn, m.w_in, L.make_epilogue(bias=m.b_in, out_f32=h, ldo=C))
        n = ws.get(tag + "n", (rows, C * Wd))
        o = ws.get(tag + "o", (rows, C * Wd))
        for a_i, a in enumerate(m.attn):
            if stream_ctx is not None:
                stream_ctx.attend(self, idx, a_i, a, h, n, o, hw, C, ws)
            elif sp:
                qkv32 = ws.get(tag + "qkv_f32", (rows, 3 * C), torch.float32)
                L.layernorm(h, a["ln_w"], a["ln_b"], n, 1e-5, C=C, pe=a["pe"], pe_rows_per_t=hw, pe_T=32, split=1)
                L.gemm(n, a["wqkv"], L.make_epilogue(out_f32=qkv32, ldo=3 * C))
                L.temporal_attn_f32(qkv32, C, o, C, hw, 8, T, T)
            else:
                qkv = ws.get(tag + "qkv", (rows, 3 * C))
                L.layernorm(h, a["ln_w"], a["ln_b"], n, 1e-5, pe=a["pe"], pe_rows_per_t=hw, pe_T=32)
                L.gemm(n, a["wqkv"], L.make_epilogue(out=qkv, ldo=3 * C))
                L.temporal_attn(qkv, 3 * C, qkv[:, C:], 3 * C, qkv[:, 2 * C:], 3 * C, o, C, hw, 8, C // 8, T, T)
            L.gemm(o, a["wo"], L.make_epilogue(bias=a["bo"], res_f32=h, out_f32=h, ldo=C))
        L.layernorm(h, m.ffn_w, m.ffn_b, n, 1e-5, C=C, split=int(sp))
        g = ws.get(tag + "geglu", (rows, 4 * C * Wd))
        L.gemm(n, m.w_geglu, self._ep(4 * C, bias=m.b_geglu, out=g, act=L.ACT_GEGLU))
        hb = ws.get(tag + "hb", (rows, C * Wd))
        L.gemm(g, m.w_ff2, self._ep(C, bias=m.b_ff2, res_f32=h, out=hb, ld32=C))
        L.gemm(hb, m.w_out, self._ep(C, bias=m.b_out, res1=x, out=out, out_relu=out_relu))
        return out

    # ------------------------------------------------------------------ DPT head
    def _fusion(self, idx, ws, T, h, w_, x0, skip, skip_relu, size, upsample=True):
        """FeatureFusionBlock (util/blocks.py:135-162) on NHWC bf16 maps of h x w_ pixels.
        x0: first input (None for refinenet4, whose only input arrives as skip/skip_relu);
        skip/skip_relu: layerN_rn output and its rectified copy.  The 1x1 out_conv is applied BEFORE
        the bilinear upsample (both linear, interpolation weights sum to 1 => same result, 1/4 FLOPs)."""
        f = self.w.fusion[idx]
        Fc = self.w.F
        Wd, Fw = self.W, self.w.F * self.W
        rows = T * h * w_
        tag = f"rf{idx}_"
        t = ws.get(tag + "t", (rows, Fw))
        if x0 is not None:
            w1, b1, w2, b2 = f.rcu[0]
            L.conv3x3(skip_relu, w1, self._ep(Fc, bias=b1, out=t, act=L.ACT_RELU), T, h, w_, Fw, Fc)
            s = ws.get(tag + "s", (rows, Fw))
            sr = ws.get(tag + "sr", (rows, Fw))
            L.conv3x3(t, w2, self._ep(Fc, bias=b2, res1=skip, res2=x0, out=s, out_relu=sr), T, h, w_, Fw, Fc)
        else:
            s, sr = skip, skip_relu
        w1, b1, w2, b2 = f.rcu[1]
        L.conv3x3(sr, w1, self._ep(Fc, bias=b1, out=t, act=L.ACT_RELU), T, h, w_, Fw, Fc)
        u = ws.get(tag + "u", (rows, Fw))
        L.conv3x3(t, w2, self._ep(Fc, bias=b2, res1=s, out=u), T, h, w_, Fw, Fc)
        v = ws.get(tag + "v", (rows, Fw))
        L.gemm(u, f.w_out, self._ep(Fc, bias=f.b_out, out=v))
        if not upsample:
            return v
        H2, W2 = size
        up = ws.get(tag + "up", (T * H2 * W2, Fw))
        sp_stride = Fc if self.fp32 else 0
        L.bilinear_nhwc(v, up, T, h, w_, H2, W2, Fc, ldx=Fw, ldo=Fw, split_in=sp_stride, split_out=sp_stride)
        return up

    def head(self, taps, gh, gw, T, ws, stream_ctx=None):
        """DPTHeadTemporal.forward (dpt_temporal.py:53-125) -> depth f32 [T, 14gh, 14gw]."""
        w = self.w
        Fc, oc, ocp, hw = w.F, w.oc, w.ocp, gh * gw
        sp, Wd = self.fp32, self.W
        rows = T * hw
        gh4, gw4 = (gh - 1) // 2 + 1, (gw - 1) // 2 + 1
        hw4 = gh4 * gw4

        def proj(i):
            wk, bk = w.proj[i]
            n_i = wk.shape[0]
            p = ws.get(f"proj{i}", (rows, n_i * Wd))
            L.gemm(taps[i], wk, self._ep(n_i, bias=bk, out=p))
            return p

        def rn(i, src, n_h, n_w, cin):
            r = ws.get(f"rn{i}", (T * n_h * n_w, Fc * Wd))
            rr = ws.get(f"rn{i}r", (T * n_h * n_w, Fc * Wd))
            L.conv3x3(src, w.rn[i], self._ep(Fc, out=r, out_relu=rr), T, n_h, n_w, cin * Wd, Fc)
            return r, rr

        # The three reassemble branches that feed refinenet 3, 2 and 1 do not depend on the refinenet-4 chain: they run
        # on side streams inside the captured graph and join where their results are consumed.  For a streamed frame
        # every launch is a fraction of a wave (vitl 4.13 -> 3.97 ms per frame); in a 32-frame window only the small
        # launches of levels 3 / 4 gain (49.37 -> 49.11 ms, profiles/r02d_ab_head_streams*.log).
        par = (self.head_streams >= 2 or (self.head_streams == 1 and stream_ctx is not None)) and L.PROFILE is None   # per-launch event timing stays serial
        joins = {}
        if par:
            main = torch.cuda.current_stream()
            fork = torch.cuda.Event()
            fork.record(main)
            if self._side is None:
                self._side = [torch.cuda.Stream(device=self.device) for _ in range(3)]

        def branch(i, fn):
            if not par:
                return fn()
            s = self._side[i]
            s.wait_event(fork)
            with torch.cuda.stream(s):
                out = fn()
                joins[i] = torch.cuda.Event()
                joins[i].record(s)
            return out

        def join(i):
            if par:
                torch.cuda.current_stream().wait_event(joins[i])

        def branch1():   # layer1: ConvTranspose x4 -> layer1_rn
            wk, bk, s_, cop = w.ct0
            l1 = ws.get("layer1", (T * 16 * hw, cop * Wd))
            L.gemm(proj(0), wk, self._ep(cop, bias=bk, out=l1, store_mode=L.ST_CONVT, ct=(s_, cop, gh, gw)), K=None if sp else oc[0])
            return rn(0, l1, 4 * gh, 4 * gw, ocp[0])

        def branch2():   # layer2: ConvTranspose x2 -> layer2_rn
            wk, bk, s_, cop = w.ct1
            l2 = ws.get("layer2", (T * 4 * hw, cop * Wd))
            L.gemm(proj(1), wk, self._ep(cop, bias=bk, out=l2, store_mode=L.ST_CONVT, ct=(s_, cop, gh, gw)), K=None if sp else oc[1])
            return rn(1, l2, 2 * gh, 2 * gw, ocp[1])

        def branch3():   # layer3: motion module 0 -> layer3_rn
            l3m = self.motion(0, proj(2), T, hw, ws, ws.get("layer3m", (rows, oc[2] * Wd)), stream_ctx=stream_ctx)
            return rn(2, l3m, gh, gw, ocp[2])

        r1, r1r = branch(0, branch1)
        r2, r2r = branch(1, branch2)
        r3, r3r = branch(2, branch3)
        p3_in = proj(3)
        col = ws.get("rs3_col", (T * hw4, 9 * oc[3] * Wd))
        L.im2col_s2(p3_in, col, T, gh, gw, oc[3] * Wd)
        l4 = ws.get("layer4", (T * hw4, oc[3] * Wd))
        L.gemm(col, w.w_rs3, self._ep(oc[3], bias=w.b_rs3, out=l4))
        l4m = self.motion(1, l4, T, hw4, ws, ws.get("layer4m", (T * hw4, oc[3] * Wd)), stream_ctx=stream_ctx)
        r4, r4r = rn(3, l4m, gh4, gw4, ocp[3])
        p4 = self._fusion(4, ws, T, gh4, gw4, None, r4, r4r, (gh, gw))
        p4m = self.motion(2, p4, T, hw, ws, ws.get("path4m", (rows, Fc * Wd)), stream_ctx=stream_ctx)
        join(2)
        p3 = self._fusion(3, ws, T, gh, gw, p4m, r3, r3r, (2 * gh, 2 * gw))
        p3m = self.motion(3, p3, T, 4 * hw, ws, ws.get("path3m", (T * 4 * hw, Fc * Wd)), stream_ctx=stream_ctx)
        join(1)
        p2 = self._fusion(2, ws, T, 2 * gh, 2 * gw, p3m, r2, r2r, (4 * gh, 4 * gw))
        F2p = w.F2p
        Ho, Wo = gh * PATCH, gw * PATCH
        depth = ws.get("depth", (T, Ho, Wo), torch.float32)
        sp_2 = F2p if sp else 0
        # Tail = refinenet1's upsample (x2) -> output_conv1 3x3 -> upsample to the input size -> 3x3(->32)+ReLU+1x1+ReLU
        # (dpt_temporal.py:96-123).  The last pair is ONE tcgen05 kernel (csrc/depth_tail.cu): the upsampled map (2.2 GB
        # per vitl window) is built tile by tile in shared memory and never touches HBM.  (Running the unfused pair per
        # frame group through one L2-sized buffer was measured slower: profiles/r02a_ab_tail_chunking_negative.log.)
        fused = (not sp) and self.tail_fused and L.depth_tail_supported(8 * gh, 8 * gw, Ho, Wo, F2p)
        join(0)
        v1 = self._fusion(1, ws, T, 4 * gh, 4 * gw, p2, r1, r1r, (8 * gh, 8 * gw))
        h8, w8 = 8 * gh, 8 * gw
        o1 = ws.get("out1", (T * h8 * w8, F2p * Wd))
        L.conv3x3(v1, w.w_oc1, self._ep(F2p, bias=w.b_oc1, out=o1), T, h8, w8, Fc * Wd, F2p)
        if fused:
            L.depth_tail(o1, w.w_oc2, w.b_oc2, w.w_last, w.b_last, depth, T, h8, w8, Ho, Wo, F2p)
        else:   # FP32-accuracy mode (split tensors) and upsampling ratios the fused kernel does not take
            up2 = ws.get("out1_up", (T * Ho * Wo, F2p * Wd))
            L.bilinear_nhwc(o1, up2, T, h8, w8, Ho, Wo, F2p, ldx=F2p * Wd, ldo=F2p * Wd, split_in=sp_2, split_out=sp_2)
            L.conv3x3(up2, w.w_oc2, L.make_epilogue(bias=w.b_oc2, out_f32=depth, w2=w.w_last, b2=w.b_last,
                                                    store_mode=L.ST_HEADDOT), T, Ho, Wo, F2p * Wd, 32)
        return depth

    # ------------------------------------------------------------------ whole window
    def forward(self, x, stream_ctx=None):
        """x f32 [T,3,H,W] on this device -> depth f32 [T,H,W] (a workspace buffer: copy before reuse).
        With `stream_ctx` (T == 1) the temporal attention reads the streaming K/V cache instead.

        The only kernel that reads `x` is the patch gather; it is launched eagerly from wherever `x` lives (no
        staging copy of the window).  Everything after it has a launch sequence that is fixed per input shape, so
        after one eager pass it is captured into a CUDA graph and replayed (one graph launch instead of ~270 kernel
        launches); `use_graph = False` or an active L.PROFILE (per-launch event timing) runs it eagerly."""
        assert x.dim() == 4 and x.shape[1] == 3 and x.dtype == torch.float32 and x.is_cuda
        T, _, H, W = x.shape
        if H % PATCH or W % PATCH:
            raise AssertionError(f"Input image size {H}x{W} is not a multiple of patch size {PATCH}")
        if T > 32:
            raise AssertionError("temporal attention supports at most 32 frames per window")
        key = ("win", T, H, W)
        ws = self._ws(key)
        with torch.cuda.device(self.device):
            x = x.contiguous()
            eager = not self.use_graph or L.PROFILE is not None
            if stream_ctx is not None:
                # Streaming step: once the cache is full (Tk == 32, from the second frame on) the launch
                # sequence is fixed -- the per-frame slot list lives in a device buffer that begin_frame()
                # refreshes -- so it is captured once per cache and replayed (per-frame latency is
                # launch-bound otherwise: ~300 launches for one frame of work).
                if eager or stream_ctx.Tk != 32:
                    return self._run(x, ws, stream_ctx)
                gkey = ("stream", T, H, W, stream_ctx.serial)
                if gkey not in self.graphs and gkey not in self._warm:
                    # first full-cache frame runs eagerly (allocates buffers); graphs of earlier streams are dropped
                    self._warm = {gkey}
                    self.graphs = {k: v for k, v in self.graphs.items() if k[0] != "stream"}
                    return self._run(x, ws, stream_ctx)
            else:
                if eager:
                    return self._run(x, ws, None)
                gkey = key
                if gkey not in self.graphs:
                    self._run(x, ws, None)               # eager pass: allocates the workspace, sets kernel attributes
            n0 = L.LAUNCHES
            prev = L.set_pdl(stream_ctx is not None)
            try:
                self._patch(x, ws)
            finally:
                L.set_pdl(prev)
            g = self.graphs.get(gkey)
            if g is None:
                torch.cuda.current_stream().synchronize()
                graph = torch.cuda.CUDAGraph()
                # torch.cuda.graph keeps ONE default capture stream per process (created on the device of the first
                # capture); a model on a second device in the same process needs a capture stream of its own device
                if self._capture_stream is None:
                    self._capture_stream = torch.cuda.Stream(device=self.device)
                with torch.cuda.graph(graph, stream=self._capture_stream):
                    out = self._after_patch(T, H, W, ws, stream_ctx)
                g = self.graphs[gkey] = (graph, out)
            g[0].replay()
            L.LAUNCHES = n0 + self.launches_per_run[key]
            return g[1]

    def _patch(self, x, ws):
        """The one consumer of the input window: 14x14 patch gather into the A operand of the patch-embed GEMM."""
        T, _, H, W = x.shape
        hw = (H // PATCH) * (W // PATCH)
        a = ws.get("patch_a", (T * hw, 592 * self.W))
        L.patch_im2col(x, a, T, H, W, 592, split=int(self.fp32))

    def _after_patch(self, T, H, W, ws, stream_ctx):
        # Programmatic dependent launch only for the latency-bound single-frame sequence of the streaming mode
        # (~300 launches of one tile each: -2 % per frame); on a full window it costs 2 % (csrc/api.cu).
        prev = L.set_pdl(stream_ctx is not None)
        try:
            taps = self.encoder(T, H, W, ws)
            return self.head(taps, H // PATCH, W // PATCH, T, ws, stream_ctx=stream_ctx)
        finally:
            L.set_pdl(prev)

    def _run(self, x, ws, stream_ctx):
        T, _, H, W = x.shape
        n0 = L.LAUNCHES
        prev = L.set_pdl(stream_ctx is not None)
        try:
            self._patch(x, ws)
        finally:
            L.set_pdl(prev)
        out = self._after_patch(T, H, W, ws, stream_ctx)
        self.launches_per_run[("win", T, H, W)] = L.LAUNCHES - n0
        return out


class StreamCache:
    """Device-side temporal cache of the streaming mode (video_depth_stream.py:55-60,134-159).

    The reference keeps, per frame, the 8 post-LayerNorm hidden states feeding the 8 temporal
    attentions and re-projects K/V of all of them every frame.  Here each of the 8 attentions owns
    a slot-addressed K/V cache [hw, cap, C] of PE-free projections; the reference's
    `frame_cache_list` becomes a list of slot numbers with identical list semantics."""

    CAP = 44  # the reference's list peaks at 43 entries (42 after the trim) + 1 slot in flight
    _SERIAL = 0  # graphs captured for a cache are keyed by this, never by id() (ids are recycled)

    def __init__(self, engine, gh, gw):
        StreamCache._SERIAL += 1
        self.serial = StreamCache._SERIAL
        w = engine.w
        dev = engine.device
        gh4, gw4 = (gh - 1) // 2 + 1, (gw - 1) // 2 + 1
        geo = [(gh * gw, w.oc[2]), (gh4 * gw4, w.oc[3]), (gh * gw, w.F), (4 * gh * gw, w.F)]
        self.fp32 = engine.fp32
        dt = torch.float32 if engine.fp32 else torch.bfloat16   # FP32-accuracy mode keeps fp32 projections
        self.k = [[torch.zeros(hw, self.CAP, C, dtype=dt, device=dev) for _ in range(2)] for hw, C in geo]
        self.v = [[torch.zeros(hw, self.CAP, C, dtype=dt, device=dev) for _ in range(2)] for hw, C in geo]
        self.slots_dev = torch.zeros(32, dtype=torch.int32, device=dev)
        self.slots_host = torch.zeros(32, dtype=torch.int32).pin_memory()
        self.Tk = 1

    def begin_frame(self, cached_slots, new_slot):
        """cached_slots: slot numbers of the list positions the new frame attends to (<= 31)."""
        lst = list(cached_slots) + [new_slot]
        self.Tk = len(lst)
        self.slots_host[: self.Tk] = torch.tensor(lst, dtype=torch.int32)
        self.slots_dev.copy_(self.slots_host, non_blocking=True)

    def attend(self, eng, idx, a_i, a, h, n, o, hw, C, ws):
        if eng.fp32:
            qkv32 = ws.get(f"mm{idx}_qkv_f32", (hw, 3 * C), torch.float32)
            L.layernorm(h, a["ln_w"], a["ln_b"], n, 1e-5, C=C, split=1)
            L.gemm(n, a["wqkv"], L.make_epilogue(out_f32=qkv32, ldo=3 * C))
            L.temporal_attn_stream_f32(qkv32, qkv32[:, C:], qkv32[:, 2 * C:], 3 * C, self.k[idx][a_i], self.v[idx][a_i],
                                       self.slots_dev, a["pe_q"], a["pe_k"], a["pe_v"], o, C, hw, 8, C // 8, self.Tk)
            return
        qkv = ws.get(f"mm{idx}_qkv", (hw, 3 * C))
        L.layernorm(h, a["ln_w"], a["ln_b"], n, 1e-5)
        L.gemm(n, a["wqkv"], L.make_epilogue(out=qkv, ldo=3 * C))
        L.temporal_attn_stream(qkv, qkv[:, C:], qkv[:, 2 * C:], 3 * C, self.k[idx][a_i], self.v[idx][a_i], self.slots_dev,
                               a["pe_q"], a["pe_k"], a["pe_v"], o, C, hw, 8, C // 8, self.Tk)

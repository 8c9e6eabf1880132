/* libvda_b200.so -- C ABI of the B200 (sm_100a) kernels behind Video Depth Anything's per-window
 * inference path.  The reference (apanner/VDA) is pure Python/PyTorch and has no FFI layer; every
 * entry point below replaces one class of torch call sites on the hot path (reference file:line in
 * each comment, paths relative to the reference root).  All pointers are DEVICE pointers unless
 * stated otherwise, `stream` is a cudaStream_t passed as void*, every function returns 0 on success
 * or a non-zero code (cudaError_t, or <0 for argument errors) and records a message readable with
 * vda_last_error().  No allocation happens inside; callers own all buffers.
 *
 * Activation layout everywhere: row-major [rows, channels] (tokens or NHWC pixels), bf16 unless a
 * parameter says f32.  Weights are bf16 [N, K] with K contiguous.
 */
#ifndef VDA_B200_H
#define VDA_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

int vda_init(int device);            /* resolve driver entry points, query SM count; idempotent */
const char* vda_last_error(void);    /* message of the last failing call on this thread */
int vda_num_sms(void);             /* SM count of the current device */
const char* vda_build_digest(void);  /* sha256 of the sources + flags this binary was built from */
/* Launch the following kernels with programmatic stream serialization (each kernel's prologue overlaps the
 * previous kernel's tail; csrc/common.cuh).  Returns the previous setting.  Default 0. */
int vda_set_pdl(int on);

enum { VDA_ACT_NONE = 0, VDA_ACT_GELU = 1, VDA_ACT_RELU = 2, VDA_ACT_GEGLU = 3 };
enum { VDA_ST_PLAIN = 0, VDA_ST_CONVT = 1, VDA_ST_QKV = 2, VDA_ST_HEADDOT = 3 };

/* Fused GEMM epilogue: v = acc + bias[col]; v = act(v); v *= scale[col];
 * v += res_f32 + res1 + res2 (each optional, addressed like the output); stores v to out (bf16),
 * out_f32 and relu(v) to out_relu (each optional). */
typedef struct vda_epilogue {
  const float* bias;    /* [N] or NULL */
  const float* scale;   /* [N] or NULL: LayerScale gamma (dinov2_layers/layer_scale.py:27-28) */
  const float* res_f32; /* fp32 residual stream (dinov2_layers/block.py:105-106) or NULL */
  const void* res1;     /* bf16 residuals (util/blocks.py:91,146) or NULL */
  const void* res2;
  void* out;            /* bf16 [rows, ldo] or NULL */
  void* out_relu;       /* bf16, relu(v): pre-activated copy for the next ResidualConvUnit */
  float* out_f32;       /* fp32 [rows, ldo] or NULL (may alias res_f32: in-place residual) */
  int ldo;              /* row pitch in elements of out/out_relu/out_f32/res* */
  int act;              /* VDA_ACT_*; GEGLU expects W rows interleaved in 16-row value/gate groups */
  int store_mode;       /* VDA_ST_* */
  int block_n;          /* 0 = auto, else 32/64/128/256 */
  /* VDA_ST_CONVT: ConvTranspose2d kernel=stride=ct_s (dpt.py:70-82); N = ct_s*ct_s*ct_cout */
  int ct_s, ct_cout, ct_h, ct_w;
  /* VDA_ST_QKV: columns [0,2D) stored plainly, V columns [2D,3D) stored transposed as
   * vt[(frame*heads+head), 64, s_pad] for the attention kernel (dinov2_layers/attention.py:51-52) */
  void* vt;
  int qkv_d, qkv_heads, qkv_s, qkv_spad;
  /* VDA_ST_HEADDOT: out_f32[row] = relu(b2 + sum_c relu(acc+bias)[c]*w2[c])  (dpt.py:118-124) */
  const float* w2;
  float b2;
  int cta_pair;         /* 0 = auto, 1 = force the single-CTA kernel, 2 = force the CTA-pair (cta_group::2) kernel */
  /* FP32-accuracy mode (infer_video_depth(..., fp32=True), video_depth.py:108): if > 0, out / out_relu /
   * res1 / res2 are "split" tensors: value x is stored as bf16 hi at [col], bf16(x - hi) at [col + split]
   * and hi again at [col + 2*split], i.e. ready to be the A operand of the next GEMM whose weights are
   * packed [Wh | Wh | Wl] along K.  `split` is the distance between the parts in elements. */
  int split;
  int ld32;             /* row pitch of out_f32 / res_f32 if different from ldo (0 = ldo); plain stores only */
} vda_epilogue;

/* C[M,N] = epi(A[M,K] * W[N,K]^T).  A bf16 with row pitch lda (elements, multiple of 8), W bf16
 * with row pitch ldw.  Replaces nn.Linear / 1x1 Conv2d / ConvTranspose2d(k=s):
 * dinov2_layers/attention.py:51,60; mlp.py:36-39; dpt.py:60-82; util/blocks.py:160;
 * motion_module/motion_module.py:115,128; motion_module/attention.py:81-83,314,382-384. */
int vda_gemm_bf16(const void* A, int lda, const void* W, int ldw, int M, int N, int K,
                  const vda_epilogue* epi, void* stream);

/* 3x3 / stride 1 / pad 1 convolution on an NHWC bf16 map as implicit GEMM (TMA-shifted taps).
 * x: [n, H, W, ldx>=C] bf16, C multiple of 64; w: [Cout, 9*C] bf16 ordered (kh, kw, c).
 * Replaces util/blocks.py:20-32,78-91 and dpt.py:117-119. */
int vda_conv3x3_nhwc(const void* x, int ldx, const void* w, int n, int H, int W, int C, int Cout,
                     const vda_epilogue* epi, void* stream);


/* Flash-style spatial attention, head_dim 64, non-causal (dinov2_layers/attention.py:53-59 / :76).
 * qk: [frames*S, 2D] bf16 (Q | K), vt: [frames*heads, 64, s_pad] bf16 (V transposed, zero padded,
 * s_pad % 8 == 0) -- both written by vda_gemm_bf16 with VDA_ST_QKV.  out: [frames*S, D] bf16.
 * fixup_flags: NULL, or frames*heads*ceil(S/128) ints that are zero on entry and on return.  The library always runs the
 * exact online softmax; the flags are only used by experiment builds (-DVDA_ATTN_EXPERIMENTS, VDA_ATTN_OPT=1: optimistic
 * softmax with the offset fixed after the first KV tile + an exact fix-up pass over flagged query tiles -- same result,
 * measured 5 % slower on B200). */
int vda_flash_attn_fwd(const void* qk, const void* vt, void* out, int frames, int S, int heads, int s_pad,
                       int* fixup_flags, void* stream);

/* LayerNorm over the last dim (dinov2_layers/block.py:84,87; dinov2.py:310; motion_module.py:175,186).
 * x: [rows, C] f32 (x_is_f32) or bf16; out bf16.  If tokens_per_frame > 0 the first `skip` tokens of
 * every frame are dropped from the output (cls token, dinov2.py:311-312).  If pe != NULL,
 * pe[(row / pe_rows_per_t) % pe_T][:] is added after normalisation (motion_module.py:207-209) and
 * out_nope (optional) receives the value before the add (the streaming cache, motion_module.py:251). */
int vda_layernorm(const void* x, int x_is_f32, const float* gamma, const float* beta, void* out, void* out_nope,
                  const float* pe, int rows, int C, float eps, int tokens_per_frame, int skip, int pe_rows_per_t,
                  int pe_T, int split, void* stream);
/* split != 0 (FP32-accuracy mode): out / out_nope rows are 3*C wide, [hi | lo | hi] (see vda_epilogue::split). */

/* GroupNorm(32 groups) on an NHWC bf16 map [n, hw, ldx>=C] -> [n, hw, C] bf16 (motion_module.py:112).
 * stats_ws: 96*n doubles of scratch (64*n fp64 sums + 32*n float2 mean/rstd). */
int vda_groupnorm_nhwc(const void* x, int ldx, const float* gamma, const float* beta, void* out, double* stats_ws,
                       int n, int hw, int C, float eps, int split, void* stream);
/* split > 0: x is a split tensor whose parts are `split` elements apart; out is [n, hw, 3*C] split, parts C apart. */
/* Allocation-free variant of vda_groupnorm_nhwc for callers that cannot allocate scratch (CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in the reference.
 * The statistics scratch is a file-scope `static __device__` array inside the library
 * (96 * VDA_GROUPNORM_STATIC_MAX_N doubles = 48 KB, held by every GPU that loads the library until exit); its device
 * address is resolved once per device with cudaGetSymbolAddress.  The call makes no cudaMalloc / cudaFree / allocating
 * library call.  n > VDA_GROUPNORM_STATIC_MAX_N is rejected (-8).  All calls on a device share the array: calls that
 * overlap on the GPU (different streams) race; serialise them or use vda_groupnorm_nhwc with your own scratch. */
#define VDA_GROUPNORM_STATIC_MAX_N 64
int vda_groupnorm_nhwc_static(const void* x, int ldx, const float* gamma, const float* beta, void* out, int n, int hw,
                              int C, float eps, int split, void* stream);

/* 14x14/stride-14 patch gather: x f32 NCHW [frames,3,H,W] -> A bf16 [frames*(H/14)*(W/14), ldk],
 * k = (c,kh,kw), zero padded to ldk (dinov2_layers/patch_embed.py:66,76-78). */
int vda_patch_im2col(const float* x, void* a, int frames, int H, int W, int ldk, int split, void* stream);
/* split != 0: A rows are 3*ldk wide, [hi | lo | hi] of the fp32 pixels. */

/* tokens[f,0]=cls+pos[0]; tokens[f,1+i]=patch[f*hw+i]+pos[1+i], all f32 (dinov2.py:218-219). */
int vda_add_cls_pos(const float* patch, const float* cls, const float* pos, float* tokens, int frames, int hw, int D,
                    void* stream);

/* Bilinear resampling with align_corners=True: NHWC bf16 (util/blocks.py:156-158,
 * dpt_temporal.py:116-118) and single-channel f32 with optional ReLU (video_depth.py:66-67,111-112). */
int vda_bilinear_nhwc(const void* x, int ldx, void* out, int ldo, int n, int h, int w, int H, int W, int C,
                      int split_in, int split_out, void* stream);
/* split_in / split_out > 0: split tensors (parts that many elements apart), interpolated in fp32. */
int vda_bilinear_f32(const float* x, float* out, int n, int h, int w, int H, int W, int relu, void* stream);

/* 3x3 stride-2 pad-1 patch gather for dpt.py:85-90: x NHWC bf16 [n,h,w,C] -> [n*ho*wo, 9*C]. */
int vda_im2col_s2(const void* x, void* a, int n, int h, int w, int C, void* stream);

/* Temporal attention core (motion_module/motion_module.py:268-308, attention.py:187-207): element
 * (t, token, c) at row t*hw + token; softmax over Tk <= 32 frames for each of Tq query frames. */
int vda_temporal_attn(const void* q, int ldq, const void* k, int ldk, const void* v, int ldv, void* out, int ldo,
                      int hw, int heads, int head_dim, int Tq, int Tk, void* stream);

/* Streaming temporal attention (video_depth_stream.py:134-153, motion_module.py:252-266): the new
 * frame (Tq = 1) attends over Tk <= 32 list positions.  kcache/vcache: [hw, cap, C] bf16 PE-free
 * projections of cached frames (the cap slots of one token are adjacent: coalesced, page-local reads);
 * slots: DEVICE int[Tk], cache slot of each list position, the last entry being the slot this call stores
 * knew/vnew into; pe_q/pe_k/pe_v: f32 [32, C] positional table projected through W_q/W_k/W_v
 * (K(n+pe) = Kn + K pe).  q/knew/vnew: [hw, ldn], out: [hw, ldo].  C = heads*head_dim <= 1024, head_dim % 8 == 0. */
int vda_temporal_attn_stream(const void* q, const void* knew, const void* vnew, int ldn, void* kcache, void* vcache,
                             const int* slots, const float* pe_q, const float* pe_k, const float* pe_v, void* out,
                             int ldo, int hw, int heads, int head_dim, int Tk, int cap, void* stream);
/* The same for the FP32-accuracy mode (infer_video_depth_one(..., fp32=True), video_depth_stream.py:77):
 * q/knew/vnew and the caches are f32, out is a split tensor (vda_epilogue::split) with parts `split` apart. */
int vda_temporal_attn_stream_f32(const float* q, const float* knew, const float* vnew, int ldn, float* kcache,
                                 float* vcache, const int* slots, const float* pe_q, const float* pe_k,
                                 const float* pe_v, void* out, int ldo, int split, int hw, int heads, int head_dim,
                                 int Tk, int cap, void* stream);

/* Fused depth tail (dpt_temporal.py:116-123, dpt.py:117-124): bilinear(align_corners=True) of the NHWC bf16 map
 * x [n,h,w,ldx>=C] to (H,W), 3x3 conv C->32 (w3 bf16 [32, 9*C], k=(kh,kw,c); b3) + ReLU, 1x1 conv 32->1 (w1, b1) + ReLU
 * -> out f32 [n,H,W], in one tcgen05 kernel that never materialises the upsampled map.  C in {64, 128}; returns -7 if
 * the upsampling ratio needs more than 8 x 12 source pixels per 10 x 18 patch (callers then use vda_bilinear_nhwc +
 * vda_conv3x3_nhwc). */
int vda_depth_tail(const void* x, int ldx, const void* w3, const float* b3, const float* w1, float b1, float* out,
                   int n, int h, int w, int H, int W, int C, void* stream);

/* Depth visualisation (utils/dc_utils.py:72-86): minmax2 = {min, max} of x as two fp32 words (any sign);
 * colorize: q = ((d - min) / (max - min) * 255).astype(uint8) evaluated with numpy's three fp32 roundings and
 * truncating cast (bit-exact), out = lut_rgb[q] (3 bytes per pixel) or q if grey. */
int vda_minmax_f32(const float* x, long long n, void* minmax2, void* stream);
int vda_depth_colorize(const float* depth, long long n, const void* minmax2, const void* lut_rgb, void* out, int grey,
                       void* stream);
/* Near/far depth normalisation of the fork's job engine (external_engine/engine.py:1099-1141), per frame:
 * t = (d - min_f) / (max_f - min_f); out = near + (far-near)*t if far > near else far + (near-far)*t; constant
 * frames -> near; invert: out = far - (out - near).  fp32 roundings as numpy's (bit-exact).
 * minmax_ws: 2*frames fp32 words of scratch (receives each frame's {min, max}). */
int vda_depth_normalize(const float* depth, float* out, long long per_frame, int frames, double near_value,
                        double far_value, int invert, void* minmax_ws, void* stream);

/* Frame pre-processing (util/transform.py:51-157 via video_depth.py:77-89): frames uint8 RGB [n,H0,W0,3]
 * -> out f32 [n,3,nh,nw] = ((cv2-compatible INTER_CUBIC resize of frame/255) - mean) / std. */
int vda_preprocess_u8(const void* frames, int n, int H0, int W0, float* out, int nh, int nw, void* stream);

/* FP32-accuracy mode attention (fp32=True of infer_video_depth, video_depth.py:108); outputs are split
 * tensors (vda_epilogue::split).
 * vda_attn_f32: q | k | v (dinov2_layers/attention.py:51-59) as a split bf16 tensor [frames*S, ld] with hi at
 * column c and lo at c + in_split (in_split >= 3*heads*64; the projection GEMM writes it with vda_epilogue::split);
 * out [frames*S, ldo] with parts `split` apart.  vda_temporal_attn_f32: as vda_temporal_attn. */
int vda_attn_f32(const void* qkv, int ld, int in_split, void* out, int ldo, int split, int frames, int S, int heads,
                 void* stream);
int vda_temporal_attn_f32(const float* q, const float* k, const float* v, int ld, void* out, int ldo, int split, int hw,
                          int heads, int head_dim, int Tq, int Tk, void* stream);

/* Alignment (utils/util.py:40-74, video_depth.py:120-160).  sums5 = {sum p*p, sum p, n, sum p*t,
 * sum t} accumulated in FP64 (caller zeroes it); solve writes {scale, shift} (metric: {1,0}). */
int vda_lsq_accum(const float* pred, const float* target, long long n, double* sums5, void* stream);
/* the same with a per-element fp32 weight (utils/util.py:16-62 with a non-trivial mask); mask NULL = all ones */
int vda_lsq_accum_masked(const float* pred, const float* target, const float* mask, long long n, double* sums5,
                         void* stream);
/* flags: bit 0 = metric model (scale 1, shift 0: video_depth.py:135-136), bit 1 = zero sums5 afterwards */
int vda_lsq_solve(double* sums5, float* scale_shift, int flags, void* stream);
/* out[f] = pre[f]*(1-w_f) + post[f]*w_f, w_f = f/(frames-1) (utils/util.py:65-74), fp32 roundings as numpy's */
int vda_blend_frames(const float* pre, const float* post, float* out, long long per_frame, int frames, void* stream);
int vda_affine_clamp(const float* src, float* dst, long long n, const float* scale_shift, void* stream);
int vda_affine_clamp_blend(float* pre, const float* post, long long per_frame, int frames, const float* scale_shift,
                           void* stream);

#ifdef __cplusplus
}
#endif
#endif

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
int vda_num_sms(void);

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

#ifdef __cplusplus
}
#endif
#endif

// Parameter block of the tcgen05 GEMM / implicit-GEMM kernel (host + device).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace vda {

// Shared-memory boxes per column slice of the TMA epilogues (gemm_epilogue.cuh), 8 KB each, two slices.
// fp32 reduce-add flavour (epi_tma 1): 2 boxes -- the issuing thread confirms that the previous box has been read
// (cp.async.bulk.wait_group.read 0) BEFORE it joins the barrier of the current box, so every thread that passes that
// barrier may overwrite the previous box; one store is still in flight while the next box is computed.  32 KB instead
// of 48 KB leave the 256-wide CTA-pair tile a fifth pipeline stage: proj 99 -> 92 us, fc2 298 -> 279 us.
// bf16 store flavours (2, 3, 4): 3 boxes with the wait after the commit; with 2 they were 2-3 % slower (fc1, K = 256).
#ifdef __CUDACC__
#define VDA_HD __host__ __device__
#else
#define VDA_HD
#endif
// cooperative flavour (5: bf16 residuals / ReLU'd second output): 2 slots x {res1 | out, res2 | out_relu} per slice.
VDA_HD constexpr int epi_boxes(int epi_tma) { return epi_tma == 1 ? 2 : (epi_tma == 5 ? 4 : 3); }
VDA_HD constexpr int epi_box_area_bytes(int epi_tma) { return 2 * epi_boxes(epi_tma) * 8 * 1024; }

// A-operand addressing
enum : int { A_PLAIN = 0, A_CONV3X3 = 1 };
// output-row addressing
enum : int {
  ST_PLAIN = 0,    // out[row*ldo + col]
  ST_CONVT = 1,    // ConvTranspose k=s: col=(kh,kw,co) scattered to the s-times larger NHWC map
  ST_QKV = 2,      // plain store + V columns also written transposed per (frame, head)
  ST_HEADDOT = 3,  // relu(acc+bias) . w2 + b2 -> relu -> out_f32[row]   (N <= BLOCK_N)
};
enum : int { ACT_NONE = 0, ACT_GELU = 1, ACT_RELU = 2, ACT_GEGLU = 3 };

struct alignas(64) GemmParams {
  CUtensorMap tmA;  // 2D [M,K] or 4D [N,H,W,C] bf16, box 64 x 128 rows, SWIZZLE_128B
  CUtensorMap tmH;  // halo mode: 4D NHWC view of x, box 64c x 10 x 18 x 1 (tile 8x16 pixels + 1-pixel halo)
  int halo;         // A_CONV3X3 only: one halo patch per channel block, 9 tap-shifted descriptors (gemm.cu)
  CUtensorMap tmB;  // 2D [N,K] bf16, box 64 x BLOCK_N
  CUtensorMap tmD;  // epi_tma: 2D [N,M] f32 view of out_f32, box 16 x 128 rows, SWIZZLE_64B
  int epi_tma;      // epilogue flavour (gemm_epilogue.cuh): 1 fp32 reduce-add, 2 bf16 box store, 3 ConvT scatter, 4 GEGLU, 5 cooperative boxes with bf16 residuals
  int M, N, K;      // GEMM extents (conv: M = images*H*W, K = 9*C)
  int num_k_blocks;
  int m_tiles, n_tiles;
  int a_mode;
  int f32_store;  // epi_tma == 1 only: store the fp32 box instead of reduce-adding it (no residual)
  int ld32;    // row pitch of out_f32 / res_f32 when it differs from ldo (0 = ldo)
  int split;   // > 0: bf16 outputs / residuals are split tensors [hi | lo | hi], parts `split` elements apart
  int stages;  // smem ring depth used by this launch (<= compiled maximum); freed stages hold the residual ring
  int sched;  // tile order: 0 = auto, 1 = round-robin, 2 = contiguous run per CTA (see gemm.cu)
  // Tail split-K (2-CTA kernel, reduce-add epilogue only): the tiles of the last, partial wave
  // [tail_begin, tiles) are each cut into `ksplit` K ranges handled by different CTA pairs, whose
  // partial sums meet in the fp32 residual through the TMA reduce-add.  ksplit <= 1 = off.
  int ksplit, tail_begin;
  // conv geometry (a_mode == A_CONV3X3): stride 1, pad 1, output map = input map
  int H, W, C;
  int tw_log2;           // spatial tile is TH x TW pixels with TW = 1<<tw_log2, TH = 128/TW
  int tiles_w, tiles_h;  // tiles per image
  // epilogue: v = acc + bias[col]; v = act(v); v *= scale[col]; v += residuals; store
  int act;
  int store_mode;
  const float* bias;      // [N] or null
  const float* scale;     // [N] or null (LayerScale gamma)
  const float* res_f32;   // fp32 residual (same addressing as out) or null
  const __nv_bfloat16* res1;  // bf16 residuals or null
  const __nv_bfloat16* res2;
  __nv_bfloat16* out;         // bf16 output or null
  __nv_bfloat16* out_relu;    // second bf16 output holding relu(v), or null
  float* out_f32;             // fp32 output or null
  int ldo;                    // row pitch (elements) of out/out_relu/out_f32/res*
  int n_out;                  // number of stored columns (N, or N/2 for GEGLU)
  // ST_CONVT
  int ct_s, ct_cout, ct_h, ct_w;  // stride(=kernel), padded Cout per tap, input map h, w
  // ST_QKV
  __nv_bfloat16* vt;  // [frames*heads, 64, s_pad]
  int qkv_d, qkv_heads, qkv_s, qkv_spad;
  // ST_HEADDOT
  const float* w2;
  float b2;
};

}  // namespace vda

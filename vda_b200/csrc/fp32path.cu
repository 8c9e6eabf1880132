// FP32-accuracy mode (infer_video_depth(..., fp32=True), video_depth.py:108; north star: rtol 1e-3).
//
// In this mode every activation that feeds a tensor-core contraction is a "split" tensor
// [hi | lo | hi] of bf16 parts (include/vda_b200.h, vda_epilogue::split) and the weights are packed
// [Wh | Wh | Wl] along K, so the tcgen05 GEMM / implicit-GEMM kernels compute hi.Wh + lo.Wh + hi.Wl
// with fp32 accumulation: a 3-term bf16 product whose relative error is ~2^-16 instead of 2^-8.
// The two attention contractions have no weights to pre-split and tiny reduction dimensions, so they
// run here as plain fp32 SIMT kernels on the fp32 q/k/v that the projection GEMM wrote:
//   vda_attn_f32          : spatial attention of the ViT blocks (dinov2_layers/attention.py:49-62)
//   vda_temporal_attn_f32 : motion-module attention over the T <= 32 frames (motion_module/attention.py:187-207)
// Both write their result as a split tensor, ready to be the A operand of the output projection.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"
#include "api_internal.h"
#include "../../include/vda_b200.h"

namespace vda {

__device__ __forceinline__ void st_split2(bf16* p, float a, float b, int split) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  *reinterpret_cast<__nv_bfloat162*>(p) = h;
  if (split) {
    const float2 f = __bfloat1622float2(h);
    *reinterpret_cast<__nv_bfloat162*>(p + split) = __floats2bfloat162_rn(a - f.x, b - f.y);
    *reinterpret_cast<__nv_bfloat162*>(p + 2 * split) = h;
  }
}

// ---------------------------------------------------------------------------------------------
// Spatial attention, head_dim 64, on warp-level tensor cores with split operands.
// Flash-attention layout: CTA = 4 warps = 64 query rows of one (frame, head), KV tiles of 64 keys.
// Every fp32 operand is split on the fly into bf16 hi + lo and every product is evaluated as
// hi.hi + lo.hi + hi.lo (QK^T additionally lo.lo) with fp32 accumulation in mma.sync.m16n8k16:
//   * Q (pre-scaled by d^-1/2 log2 e) is split once into A fragments held in registers;
//   * K / V tiles are split while they are staged global -> shared (rows padded for ldmatrix);
//   * S, the online softmax and O live in the fp32 accumulator registers; P is split into A fragments
//     directly from the accumulator layout.
// 3x (4x) the bf16 MMA work of the BF16 kernel, but no fp32 FMA inner loop: ~5x faster than the SIMT kernel.
__device__ __forceinline__ void f32_ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void f32_ldsm_x4_t(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void f32_mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// (a, b) -> packed bf16 hi pair and packed bf16 lo pair
__device__ __forceinline__ void split_pair(float a, float b, uint32_t& hi, uint32_t& lo) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  const float2 f = __bfloat1622float2(h);
  const __nv_bfloat162 l = __floats2bfloat162_rn(a - f.x, b - f.y);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = *reinterpret_cast<const uint32_t*>(&l);
}

constexpr int kAmRows = 64;          // query rows per CTA
constexpr int kAmKeys = 64;          // keys per KV tile
constexpr int kAmPitch = 64 * 2 + 16;  // bytes per staged bf16 row

// ---------------------------------------------------------------------------------------------
// Same contraction for q/k/v that the projection GEMM already wrote as split tensors (hi at column c,
// lo at c + in_split): nothing is converted in the loop, K/V tiles (hi and lo) stream global -> shared with
// cp.async into a 2-stage ring while the previous tile is in the MMAs, the d^-1/2 log2 e scale is
// folded into the softmax FFMA, and every product is 3 terms (hi.hi + lo.hi + hi.lo).  3 CTAs per SM.
__device__ __forceinline__ void f32_cp16(uint32_t dst, const void* src, int bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
constexpr int kApMat = kAmKeys * kAmPitch;   // one staged matrix (64 keys x 64 bf16, padded rows)
constexpr int kApStage = 4 * kApMat;         // K hi, K lo, V hi, V lo
constexpr int kApSmem = 2 * kApStage;

__global__ void __launch_bounds__(128, 3) attn_split_mma_kernel(const bf16* __restrict__ qkv, int ld, int in_split,
                                                                bf16* __restrict__ out, int ldo, int split, int S, int D,
                                                                float scale_log2) {
  extern __shared__ __align__(16) uint8_t sm_dyn[];
  const uint32_t sbase = smem_u32(sm_dyn);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int head = blockIdx.y, frame = blockIdx.z;
  const int g = lane >> 2, q4 = lane & 3;
  const bf16* base = qkv + static_cast<long long>(frame) * S * ld + head * 64;
  const int row_a = blockIdx.x * kAmRows + warp * 16 + g, row_b = row_a + 8;

  auto load_tile = [&](int j0, int stage) {
    const uint32_t st = sbase + stage * kApStage;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int idx = i * 128 + threadIdx.x;
      const int mat = idx >> 9, j = (idx >> 3) & 63, c = idx & 7;   // mat: 0 Kh, 1 Kl, 2 Vh, 3 Vl
      const bool ok = j0 + j < S;
      const bf16* src = base + static_cast<long long>(ok ? j0 + j : 0) * ld + ((mat & 2) ? 2 * D : D) +
                        ((mat & 1) ? in_split : 0) + c * 8;
      f32_cp16(st + mat * kApMat + j * kAmPitch + c * 16, src, ok ? 16 : 0);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  load_tile(0, 0);

  uint32_t qh[4][4], ql[4][4];
  {
    const bf16* qa = base + static_cast<long long>(min(row_a, S - 1)) * ld;
    const bf16* qb = base + static_cast<long long>(min(row_b, S - 1)) * ld;
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      const int c = kk * 16 + q4 * 2;
      qh[kk][0] = __ldg(reinterpret_cast<const uint32_t*>(qa + c)), ql[kk][0] = __ldg(reinterpret_cast<const uint32_t*>(qa + in_split + c));
      qh[kk][1] = __ldg(reinterpret_cast<const uint32_t*>(qb + c)), ql[kk][1] = __ldg(reinterpret_cast<const uint32_t*>(qb + in_split + c));
      qh[kk][2] = __ldg(reinterpret_cast<const uint32_t*>(qa + c + 8)), ql[kk][2] = __ldg(reinterpret_cast<const uint32_t*>(qa + in_split + c + 8));
      qh[kk][3] = __ldg(reinterpret_cast<const uint32_t*>(qb + c + 8)), ql[kk][3] = __ldg(reinterpret_cast<const uint32_t*>(qb + in_split + c + 8));
    }
  }
  float o[8][4];
#pragma unroll
  for (int nt = 0; nt < 8; ++nt)
#pragma unroll
    for (int i = 0; i < 4; ++i) o[nt][i] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};   // m_run in the scaled log2 domain

  const int ntiles = (S + kAmKeys - 1) / kAmKeys;
  for (int t = 0; t < ntiles; ++t) {
    const int j0 = t * kAmKeys;
    // one barrier per tile: my copies of tile t have landed, and after the barrier so have everyone's --
    // and every warp is done with tile t-1, whose stage the prefetch of tile t+1 now overwrites
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (t + 1 < ntiles) load_tile(j0 + kAmKeys, (t + 1) & 1);
    const uint32_t sKh = sbase + (t & 1) * kApStage, sKl = sKh + kApMat, sVh = sKl + kApMat, sVl = sVh + kApMat;

    float s[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
      for (int i = 0; i < 4; ++i) s[nt][i] = 0.f;
    // (the three terms of one accumulator are dependent MMAs: issue them term-major over four independent
    // accumulators so that consecutive instructions never wait on each other)
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int n2 = 0; n2 < 2; ++n2) {
        uint32_t bh[2][4], bl[2][4];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int np = 2 * n2 + e;
          const uint32_t off = (np * 16 + (lane & 7) + ((lane >> 4) << 3)) * kAmPitch + (kk * 2 + ((lane >> 3) & 1)) * 16;
          f32_ldsm_x4(sKh + off, bh[e]);
          f32_ldsm_x4(sKl + off, bl[e]);
        }
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
          for (int u = 0; u < 2; ++u) f32_mma(s[4 * n2 + 2 * e + u], qh[kk], bh[e][2 * u], bh[e][2 * u + 1]);
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
          for (int u = 0; u < 2; ++u) f32_mma(s[4 * n2 + 2 * e + u], ql[kk], bh[e][2 * u], bh[e][2 * u + 1]);
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
          for (int u = 0; u < 2; ++u) f32_mma(s[4 * n2 + 2 * e + u], qh[kk], bl[e][2 * u], bl[e][2 * u + 1]);
      }
    }
    const bool ragged = j0 + kAmKeys > S;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float m = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          if (ragged && j0 + nt * 8 + q4 * 2 + i >= S) s[nt][2 * h + i] = -INFINITY;
          m = fmaxf(m, s[nt][2 * h + i]);
        }
      m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
      m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
      m = fmaxf(m_run[h], m * scale_log2);
      const float alpha = exp2f(m_run[h] - m);
      m_run[h] = m;
      float l = 0.f;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const float e = exp2f(fmaf(s[nt][2 * h + i], scale_log2, -m));
          s[nt][2 * h + i] = e;
          l += e;
        }
      l += __shfl_xor_sync(0xffffffffu, l, 1);
      l += __shfl_xor_sync(0xffffffffu, l, 2);
      l_run[h] = l_run[h] * alpha + l;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) o[nt][2 * h] *= alpha, o[nt][2 * h + 1] *= alpha;
    }
#pragma unroll
    for (int k2 = 0; k2 < 4; ++k2) {
      uint32_t ph[4], pl[4];
      split_pair(s[2 * k2][0], s[2 * k2][1], ph[0], pl[0]);
      split_pair(s[2 * k2][2], s[2 * k2][3], ph[1], pl[1]);
      split_pair(s[2 * k2 + 1][0], s[2 * k2 + 1][1], ph[2], pl[2]);
      split_pair(s[2 * k2 + 1][2], s[2 * k2 + 1][3], ph[3], pl[3]);
#pragma unroll
      for (int n2 = 0; n2 < 2; ++n2) {
        uint32_t vh[2][4], vl[2][4];
        const int mi = lane >> 3;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int np = 2 * n2 + e;
          const uint32_t off = (k2 * 16 + (mi & 1) * 8 + (lane & 7)) * kAmPitch + (np * 16 + (mi >> 1) * 8) * 2;
          f32_ldsm_x4_t(sVh + off, vh[e]);
          f32_ldsm_x4_t(sVl + off, vl[e]);
        }
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
          for (int u = 0; u < 2; ++u) f32_mma(o[4 * n2 + 2 * e + u], ph, vh[e][2 * u], vh[e][2 * u + 1]);
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
          for (int u = 0; u < 2; ++u) f32_mma(o[4 * n2 + 2 * e + u], pl, vh[e][2 * u], vh[e][2 * u + 1]);
#pragma unroll
        for (int e = 0; e < 2; ++e)
#pragma unroll
          for (int u = 0; u < 2; ++u) f32_mma(o[4 * n2 + 2 * e + u], ph, vl[e][2 * u], vl[e][2 * u + 1]);
      }
    }
  }
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int row = h ? row_b : row_a;
    if (row < S) {
      const float inv = 1.f / l_run[h];
      bf16* dst = out + (static_cast<long long>(frame) * S + row) * ldo + head * 64;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) st_split2(dst + nt * 8 + q4 * 2, o[nt][2 * h] * inv, o[nt][2 * h + 1] * inv, split);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Temporal attention: one warp per (token, head); K and V rows of the Tk frames staged in shared
// memory as fp32, lane t owns query frame t (scores and softmax in its registers).
// Element (t, token d, channel c) of q/k/v/out lives at row t*hw + d (motion_module.py:250).
constexpr int kTfWarps = 2;

__global__ void __launch_bounds__(kTfWarps * 32) temporal_attn_f32_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                                          const float* __restrict__ v, int ld, bf16* __restrict__ out,
                                                                          int ldo, int split, int hw, int heads, int hd, int Tq,
                                                                          int Tk, float scale) {
  extern __shared__ __align__(16) float tf_sm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int unit = blockIdx.x * kTfWarps + warp;
  if (unit >= hw * heads) return;
  const int d = unit / heads, head = unit - d * heads;
  float* sK = tf_sm + warp * 2 * 32 * hd;
  float* sV = sK + 32 * hd;
  const int c4 = hd >> 2;
  for (int i = lane; i < Tk * c4; i += 32) {
    const int t = i / c4, c = i - t * c4;
    const long long row = static_cast<long long>(t) * hw + d;
    reinterpret_cast<float4*>(sK + t * hd)[c] = __ldg(reinterpret_cast<const float4*>(k + row * ld + head * hd) + c);
    reinterpret_cast<float4*>(sV + t * hd)[c] = __ldg(reinterpret_cast<const float4*>(v + row * ld + head * hd) + c);
  }
  __syncwarp();
  const int t = lane < Tq ? lane : 0;
  const float4* qr = reinterpret_cast<const float4*>(q + (static_cast<long long>(t) * hw + d) * ld + head * hd);
  float s[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) s[j] = 0.f;
  for (int c = 0; c < c4; ++c) {
    const float4 qq = __ldg(qr + c);
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (j < Tk) {
        const float4 kk = reinterpret_cast<const float4*>(sK + j * hd)[c];
        s[j] += qq.x * kk.x + qq.y * kk.y + qq.z * kk.z + qq.w * kk.w;
      }
    }
  }
  float m = -INFINITY;
#pragma unroll
  for (int j = 0; j < 32; ++j)
    if (j < Tk) m = fmaxf(m, s[j] * scale);
  float l = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    s[j] = (j < Tk) ? expf(s[j] * scale - m) : 0.f;
    l += s[j];
  }
  const float inv = 1.f / l;
  bf16* dst = out + (static_cast<long long>(t) * hw + d) * ldo + head * hd;
  for (int c = 0; c < c4; ++c) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (j < Tk) {
        const float4 vv = reinterpret_cast<const float4*>(sV + j * hd)[c];
        acc.x += s[j] * vv.x, acc.y += s[j] * vv.y, acc.z += s[j] * vv.z, acc.w += s[j] * vv.w;
      }
    }
    if (lane < Tq) {
      st_split2(dst + 4 * c, acc.x * inv, acc.y * inv, split);
      st_split2(dst + 4 * c + 2, acc.z * inv, acc.w * inv, split);
    }
  }
}

}  // namespace vda

using namespace vda;

extern "C" int vda_attn_f32(const void* qkv, int ld, int in_split, void* out, int ldo, int split, int frames, int S,
                            int heads, void* stream) {
  const int D = heads * 64;
  if (!qkv || !out || frames <= 0 || S <= 0 || heads <= 0 || ld % 8 || ld < 3 * D || in_split % 8 ||
      ldo % 8 || split % 8)
    return set_error(-1, "vda_attn_f32: bad arguments (head_dim 64, ld%%8, in_split%%8, ldo%%8, split%%8)");
  const float scale_log2 = 0.125f * 1.4426950408889634f;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (in_split <= 0) return set_error(-1, "vda_attn_f32: q/k/v must arrive as the split tensor the projection GEMM writes (in_split > 0)");
  if (ld < in_split + 3 * D) return set_error(-1, "vda_attn_f32: ld < in_split + 3*heads*64");
  static DeviceOnce once;
  int& configured = once.slot();
  if (!configured) {
    int rc = check_cuda(cudaFuncSetAttribute(attn_split_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kApSmem),
                        "attn_split smem attribute");
    if (rc) return rc;
    configured = 1;
  }
  const dim3 grid((S + kAmRows - 1) / kAmRows, heads, frames);
  attn_split_mma_kernel<<<grid, 128, kApSmem, st>>>(static_cast<const bf16*>(qkv), ld, in_split, static_cast<bf16*>(out), ldo,
                                                    split, S, D, scale_log2);
  return check_cuda(cudaGetLastError(), "attn_f32 launch");
}

extern "C" int vda_temporal_attn_f32(const float* q, const float* k, const float* v, int ld, void* out, int ldo, int split,
                                     int hw, int heads, int head_dim, int Tq, int Tk, void* stream) {
  if (!q || !k || !v || !out || hw <= 0 || heads <= 0 || head_dim % 4 || head_dim > 128 || Tq < 1 || Tq > 32 || Tk < 1 ||
      Tk > 32 || ld % 4 || ldo % 8 || split % 8)
    return set_error(-1, "vda_temporal_attn_f32: bad arguments (head_dim%%4, T<=32)");
  const int smem = kTfWarps * 2 * 32 * head_dim * 4;
  static DeviceOnce once;
  int& configured = once.slot();
  if (smem > configured) {
    int rc = check_cuda(cudaFuncSetAttribute(temporal_attn_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             kTfWarps * 2 * 32 * 128 * 4),
                        "temporal f32 smem attribute");
    if (rc) return rc;
    configured = kTfWarps * 2 * 32 * 128 * 4;
  }
  const int units = hw * heads;
  temporal_attn_f32_kernel<<<(units + kTfWarps - 1) / kTfWarps, kTfWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(
      q, k, v, ld, static_cast<bf16*>(out), ldo, split, hw, heads, head_dim, Tq, Tk,
      1.0f / sqrtf(static_cast<float>(head_dim)));
  return check_cuda(cudaGetLastError(), "temporal_attn_f32 launch");
}

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
// Spatial attention, head_dim 64.  Block = 128 threads = 128 query rows of one (frame, head);
// thread r keeps its query (pre-scaled by d^-1/2 * log2 e) and its output row in registers and
// walks the keys in chunks of 32 staged in shared memory (broadcast reads): online softmax with one
// rescale per chunk.
constexpr int kAfRows = 128;
constexpr int kAfChunk = 32;

__global__ void __launch_bounds__(kAfRows) attn_f32_kernel(const float* __restrict__ qkv, int ld, bf16* __restrict__ out,
                                                           int ldo, int split, int S, int D, float scale_log2) {
  __shared__ __align__(16) float sK[kAfChunk][64];
  __shared__ __align__(16) float sV[kAfChunk][64];
  const int head = blockIdx.y, frame = blockIdx.z;
  const int r = blockIdx.x * kAfRows + threadIdx.x;
  const float* base = qkv + static_cast<long long>(frame) * S * ld + head * 64;
  float q[64], o[64];
  {
    const float4* qr = reinterpret_cast<const float4*>(base + static_cast<long long>(r < S ? r : S - 1) * ld);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const float4 t = __ldg(qr + i);
      q[4 * i] = t.x * scale_log2, q[4 * i + 1] = t.y * scale_log2, q[4 * i + 2] = t.z * scale_log2,
             q[4 * i + 3] = t.w * scale_log2;
    }
  }
#pragma unroll
  for (int i = 0; i < 64; ++i) o[i] = 0.f;
  float m_run = -INFINITY, l_run = 0.f;

  for (int j0 = 0; j0 < S; j0 += kAfChunk) {
    __syncthreads();
    // 32 keys x 64 floats = 512 float4 per matrix, 4 per thread, rows contiguous (256 B)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = i * kAfRows + threadIdx.x;
      const int j = idx >> 4, c = idx & 15;
      float4 kk = make_float4(0.f, 0.f, 0.f, 0.f), vv = kk;
      if (j0 + j < S) {
        const float* row = base + static_cast<long long>(j0 + j) * ld;
        kk = __ldg(reinterpret_cast<const float4*>(row + D) + c);
        vv = __ldg(reinterpret_cast<const float4*>(row + 2 * D) + c);
      }
      reinterpret_cast<float4*>(&sK[j][0])[c] = kk;
      reinterpret_cast<float4*>(&sV[j][0])[c] = vv;
    }
    __syncthreads();
    float s[kAfChunk];
#pragma unroll
    for (int j = 0; j < kAfChunk; ++j) {
      float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const float4 kk = reinterpret_cast<const float4*>(&sK[j][0])[c];
        a0 = fmaf(q[4 * c], kk.x, a0), a1 = fmaf(q[4 * c + 1], kk.y, a1);
        a2 = fmaf(q[4 * c + 2], kk.z, a2), a3 = fmaf(q[4 * c + 3], kk.w, a3);
      }
      s[j] = (j0 + j < S) ? (a0 + a1) + (a2 + a3) : -INFINITY;
    }
    float m_new = m_run;
#pragma unroll
    for (int j = 0; j < kAfChunk; ++j) m_new = fmaxf(m_new, s[j]);
    const float alpha = exp2f(m_run - m_new);  // first chunk: exp2(-inf) = 0
    m_run = m_new;
    l_run *= alpha;
#pragma unroll
    for (int i = 0; i < 64; ++i) o[i] *= alpha;
#pragma unroll
    for (int j = 0; j < kAfChunk; ++j) {
      const float pj = exp2f(s[j] - m_new);
      l_run += pj;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const float4 vv = reinterpret_cast<const float4*>(&sV[j][0])[c];
        o[4 * c] = fmaf(pj, vv.x, o[4 * c]), o[4 * c + 1] = fmaf(pj, vv.y, o[4 * c + 1]);
        o[4 * c + 2] = fmaf(pj, vv.z, o[4 * c + 2]), o[4 * c + 3] = fmaf(pj, vv.w, o[4 * c + 3]);
      }
    }
  }
  if (r < S) {
    const float inv = 1.f / l_run;
    bf16* dst = out + (static_cast<long long>(frame) * S + r) * ldo + head * 64;
#pragma unroll
    for (int i = 0; i < 64; i += 2) st_split2(dst + i, o[i] * inv, o[i + 1] * inv, split);
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

extern "C" int vda_attn_f32(const float* qkv, int ld, void* out, int ldo, int split, int frames, int S, int heads,
                            void* stream) {
  const int D = heads * 64;
  if (!qkv || !out || frames <= 0 || S <= 0 || heads <= 0 || ld % 4 || ld < 3 * D || ldo % 8 || split % 8)
    return set_error(-1, "vda_attn_f32: bad arguments (head_dim 64, ld%%4, ldo%%8, split%%8)");
  const dim3 grid((S + kAfRows - 1) / kAfRows, heads, frames);
  attn_f32_kernel<<<grid, kAfRows, 0, static_cast<cudaStream_t>(stream)>>>(
      qkv, ld, static_cast<bf16*>(out), ldo, split, S, D, 0.125f * 1.4426950408889634f);
  return check_cuda(cudaGetLastError(), "attn_f32 launch");
}

extern "C" int vda_temporal_attn_f32(const float* q, const float* k, const float* v, int ld, void* out, int ldo, int split,
                                     int hw, int heads, int head_dim, int Tq, int Tk, void* stream) {
  if (!q || !k || !v || !out || hw <= 0 || heads <= 0 || head_dim % 4 || head_dim > 128 || Tq < 1 || Tq > 32 || Tk < 1 ||
      Tk > 32 || ld % 4 || ldo % 8 || split % 8)
    return set_error(-1, "vda_temporal_attn_f32: bad arguments (head_dim%%4, T<=32)");
  const int smem = kTfWarps * 2 * 32 * head_dim * 4;
  static int configured = 0;
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

// Fused epilogue shared by the 1-CTA and 2-CTA tcgen05 GEMM kernels: TMEM accumulator tile
// (128 rows x BN fp32 columns, thread = row) -> bias / activation / LayerScale / residuals -> global.
//
// Latency design: the epilogue runs on 8 warps beside the TMA and MMA warps, two per scheduler, so
// every dependent global load on its path is exposed.  Therefore
//   * bias and scale of the tile's columns are staged in shared memory once per tile (broadcast LDS),
//   * the fp32 residual rows of the tile are prefetched into L2 when the tile's epilogue starts
//     (before the MMAs of that tile have even finished) and requested before the TMEM wait,
//   * accumulator chunks are read 32 columns at a time with the next tcgen05.ld in flight while the
//     current chunk is processed.
#pragma once
#include "common.cuh"
#include "gemm_params.h"

namespace vda {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;
constexpr int kEpiWarps = 8;
constexpr int kEpiThreads = kEpiWarps * 32;

struct RowMap {
  bool valid;
  long long row;  // logical output row (pixel index / token index)
};

__device__ __forceinline__ RowMap map_row(const GemmParams& p, int m_tile, int r) {
  RowMap rm;
  if (p.a_mode == A_CONV3X3) {
    const int tpi = p.tiles_w * p.tiles_h;
    const int img = m_tile / tpi;
    const int t = m_tile - img * tpi;
    const int ty = t / p.tiles_w, tx = t - ty * p.tiles_w;
    const int tw = 1 << p.tw_log2;
    const int h = ty * (kBlockM >> p.tw_log2) + (r >> p.tw_log2);
    const int w = tx * tw + (r & (tw - 1));
    rm.valid = (m_tile < p.m_tiles) && (h < p.H) && (w < p.W);
    rm.row = (static_cast<long long>(img) * p.H + h) * p.W + w;
  } else {
    rm.row = static_cast<long long>(m_tile) * kBlockM + r;
    rm.valid = rm.row < p.M;
  }
  return rm;
}

__device__ __forceinline__ void store_bf16x16(bf16* dst, const float (&v)[16]) {
  uint32_t w[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    w[i] = *reinterpret_cast<uint32_t*>(&h);
  }
  uint4* d = reinterpret_cast<uint4*>(dst);
  d[0] = make_uint4(w[0], w[1], w[2], w[3]);
  d[1] = make_uint4(w[4], w[5], w[6], w[7]);
}

__device__ __forceinline__ void add_bf16x16(const bf16* src, float (&v)[16]) {
  const uint4* s = reinterpret_cast<const uint4*>(src);
  uint4 a = __ldg(s), b = __ldg(s + 1);
  const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
    float2 f = __bfloat1622float2(h);
    v[2 * i] += f.x;
    v[2 * i + 1] += f.y;
  }
}

__device__ __forceinline__ void load_f32x16(const float* src, float (&o)[16]) {
  const float4* s = reinterpret_cast<const float4*>(src);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float4 t = __ldg(s + i);
    o[4 * i] = t.x, o[4 * i + 1] = t.y, o[4 * i + 2] = t.z, o[4 * i + 3] = t.w;
  }
}

__device__ __forceinline__ void lds_f32x16(const float* src, float (&o)[16]) {
  const float4* s = reinterpret_cast<const float4*>(src);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float4 t = s[i];
    o[4 * i] = t.x, o[4 * i + 1] = t.y, o[4 * i + 2] = t.z, o[4 * i + 3] = t.w;
  }
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ void prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// Stage bias/scale of the tile's BN columns into shared memory (called by all 256 epilogue threads,
// et = 0..255).  s_bias/s_scale: BN floats each, double-buffered by the caller per tile parity.
template <int BN>
__device__ __forceinline__ void stage_bias_scale(const GemmParams& p, int n_t, int et, float* s_bias, float* s_scale) {
  if (et < BN) {
    const int col = n_t * BN + et;
    const bool in = col < p.N;
    s_bias[et] = (p.bias && in) ? __ldg(p.bias + col) : 0.f;
    s_scale[et] = (p.scale && in) ? __ldg(p.scale + col) : 1.f;
  }
  named_bar_sync(1, kEpiThreads);
}

// One 16-column group of one row: v already holds acc + bias.
__device__ __forceinline__ void finish_group16(const GemmParams& p, const RowMap& rm, int col, float (&v)[16],
                                               const float* s_scale16, const float4* res_pre) {
  if (p.act == ACT_GELU) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = gelu_fast(v[i]);
  } else if (p.act == ACT_RELU) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
  }
  if (p.scale) {
    float sc[16];
    lds_f32x16(s_scale16, sc);
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] *= sc[i];
  }
  long long off;
  if (p.store_mode == ST_CONVT) {
    const int gg = col / p.ct_cout, co = col - gg * p.ct_cout;
    const int kh = gg / p.ct_s, kw = gg - kh * p.ct_s;
    const int hw = p.ct_h * p.ct_w;
    const int m = static_cast<int>(rm.row);
    const int n = m / hw, rem = m - n * hw;
    const int ih = rem / p.ct_w, iw = rem - ih * p.ct_w;
    const long long orow =
        (static_cast<long long>(n) * p.ct_h * p.ct_s + ih * p.ct_s + kh) * (p.ct_w * p.ct_s) + iw * p.ct_s + kw;
    off = orow * p.ldo + co;
  } else {
    off = rm.row * p.ldo + col;
  }
  if (p.store_mode == ST_QKV && col >= 2 * p.qkv_d) {
    const int m = static_cast<int>(rm.row);
    const int b = m / p.qkv_s, sidx = m - b * p.qkv_s;
    const int c2 = col - 2 * p.qkv_d;
    const int head = c2 >> 6, d0 = c2 & 63;
    bf16* dst = p.vt + (static_cast<long long>(b * p.qkv_heads + head) * 64 + d0) * p.qkv_spad + sidx;
#pragma unroll
    for (int i = 0; i < 16; ++i) dst[static_cast<long long>(i) * p.qkv_spad] = __float2bfloat16(v[i]);
    return;
  }
  if (res_pre) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 t = res_pre[i];
      v[4 * i] += t.x, v[4 * i + 1] += t.y, v[4 * i + 2] += t.z, v[4 * i + 3] += t.w;
    }
  } else if (p.res_f32) {
    float t[16];
    load_f32x16(p.res_f32 + off, t);
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += t[i];
  }
  if (p.res1) add_bf16x16(p.res1 + off, v);
  if (p.res2) add_bf16x16(p.res2 + off, v);
  if (p.out_f32) {
    float4* d = reinterpret_cast<float4*>(p.out_f32 + off);
#pragma unroll
    for (int i = 0; i < 4; ++i) d[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  }
  if (p.out) store_bf16x16(p.out + off, v);
  if (p.out_relu) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
    store_bf16x16(p.out_relu + off, v);
  }
}

// Requests the fp32 residual lines this thread will read for the tile into L2 (no registers held).
template <int BN>
__device__ __forceinline__ void epilogue_prefetch(const GemmParams& p, int m_tile, int n_t, int r, int half) {
  if (p.res_f32 == nullptr || p.store_mode != ST_PLAIN) return;
  constexpr int kPer = (BN / 2 >= 32) ? BN / 2 : 32;
  const int cbeg = half * kPer;
  if (cbeg >= BN) return;
  const RowMap rm = map_row(p, m_tile, r);
  if (!rm.valid) return;
  const int col = n_t * BN + cbeg;
  const float* src = p.res_f32 + rm.row * p.ldo + col;
#pragma unroll
  for (int i = 0; i < kPer; i += 32)
    if (col + i < p.N) prefetch_l2(src + i);
}

// t_acc: TMEM address (lane quadrant of this warp, accumulator stage, column 0 of the tile).
// s_bias / s_scale: staged bias / scale of the tile's BN columns (see stage_bias_scale).
template <int BN>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, int m_tile, int n_t, uint32_t t_acc, int r, int half,
                                              const float* s_bias, const float* s_scale) {
  const RowMap rm = map_row(p, m_tile, r);
  constexpr int kPer = (BN / 2 >= 32) ? BN / 2 : 32;
  constexpr int kChunks = kPer / 32;
  const int cbeg = half * kPer;
  if (cbeg >= BN) return;

  if (p.store_mode == ST_HEADDOT) {
    // out_f32[row] = relu(b2 + sum_c relu(acc + bias)[c] * w2[c]), N <= 32   (dpt.py:118-124)
    uint32_t raw[32];
    tmem_ld32(t_acc, raw);
    tmem_ld_wait();
    float dot = p.b2;
#pragma unroll
    for (int g = 0; g < 2; ++g) {
      if (16 * g < p.N) {
        float w[16];
        load_f32x16(p.w2 + 16 * g, w);
#pragma unroll
        for (int i = 0; i < 16; ++i) dot += fmaxf(__uint_as_float(raw[16 * g + i]) + s_bias[16 * g + i], 0.f) * w[i];
      }
    }
    if (rm.valid) p.out_f32[rm.row] = fmaxf(dot, 0.f);
    return;
  }

  // One 16-column group per iteration, deliberately NOT unrolled: the generic epilogue body is a
  // few hundred instructions and an unrolled tile loop overflowed the instruction cache (ncu:
  // "no_instruction" stalls, 134 KB of SASS), which made the epilogue, not the MMA, set the pace.
  const bool pre = p.res_f32 != nullptr && p.store_mode == ST_PLAIN && rm.valid;
  constexpr int kStep = 16;
  const int step = (p.act == ACT_GEGLU) ? 32 : kStep;
#pragma unroll 1
  for (int c0 = cbeg; c0 < cbeg + kPer && c0 < BN; c0 += step) {
    const int col = n_t * BN + c0;
    uint32_t raw[16], gate[16];
    __syncwarp();
    tmem_ld16(t_acc + c0, raw);
    if (p.act == ACT_GEGLU) tmem_ld16(t_acc + c0 + 16, gate);
    float4 rpre[4];
    if (pre && col < p.N) {
      const float4* src = reinterpret_cast<const float4*>(p.res_f32 + rm.row * p.ldo + col);
#pragma unroll
      for (int i = 0; i < 4; ++i) rpre[i] = __ldcs(src + i);
    }
    tmem_ld_wait();
    if (!rm.valid || col >= p.N) continue;
    float v[16], b[16];
    lds_f32x16(s_bias + c0, b);
    if (p.act == ACT_GEGLU) {
      // physical columns [32j, 32j+16) = value, [32j+16, 32j+32) = gate -> output columns [16j, 16j+16)
      float bg[16];
      lds_f32x16(s_bias + c0 + 16, bg);
#pragma unroll
      for (int i = 0; i < 16; ++i)
        v[i] = (__uint_as_float(raw[i]) + b[i]) * gelu_fast(__uint_as_float(gate[i]) + bg[i]);
      store_bf16x16(p.out + rm.row * p.ldo + (col >> 1), v);
      continue;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(raw[i]) + b[i];
    finish_group16(p, rm, col, v, s_scale + c0, pre ? rpre : nullptr);
  }
}

}  // namespace vda

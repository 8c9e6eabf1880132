// Bandwidth-bound kernels of the window path: LayerNorm, GroupNorm (NHWC), patch im2col,
// cls/pos-embed assembly, bilinear (align_corners=True) resampling, stride-2 im2col.
// All are coalesced along the channel (innermost) dimension with 8/16-byte vector accesses.
#include <string.h>

#include "common.cuh"
#include "api_internal.h"
#include "../../include/vda_b200.h"

namespace vda {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------ split tensors (FP32-accuracy mode)
// value = hi + lo, stored [hi | lo | hi] with `split` elements between the parts (include/vda_b200.h,
// vda_epilogue::split).  split == 0: ordinary bf16 tensor.
__device__ __forceinline__ float2 ld_pair(const bf16* p, int split) {
  float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
  if (split) {
    const float2 l = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p + split));
    t.x += l.x, t.y += l.y;
  }
  return t;
}
__device__ __forceinline__ void st_pair(bf16* p, float a, float b, int split) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  *reinterpret_cast<__nv_bfloat162*>(p) = h;
  if (split) {
    const float2 f = __bfloat1622float2(h);
    *reinterpret_cast<__nv_bfloat162*>(p + split) = __floats2bfloat162_rn(a - f.x, b - f.y);
    *reinterpret_cast<__nv_bfloat162*>(p + 2 * split) = h;
  }
}

// ------------------------------------------------------------------ LayerNorm
// One warp per row; the row lives in registers (C <= 64*MAXV).  Two-pass (mean, then centred variance) like torch's CPU/GPU kernels.
// Output row remap: input rows are [frames, S] tokens; the first `skip` tokens of every frame are
// dropped (cls token for the DPT taps, dinov2.py:311-312).  Optional positional-encoding add
// (motion_module.py:207-209) after normalisation: pe[(row / pe_rows_per_t) % pe_T][c].
// 4 blocks of 8 warps per SM (64 registers, a 16-byte spill) instead of the 3 that 70 registers allowed: C = 1024 rows
// 54 -> 49 us (5.5 TB/s = 83 % of the measured HBM peak); 5 blocks spill more and are slower (54.5 us).
#ifndef VDA_LN_MINBLOCKS
#define VDA_LN_MINBLOCKS 4
#endif
template <typename TIn, int MAXV>
__global__ void __launch_bounds__(256, VDA_LN_MINBLOCKS) layernorm_kernel(const TIn* __restrict__ x, const float* __restrict__ gamma,
                                 const float* __restrict__ beta, bf16* __restrict__ out, bf16* __restrict__ out_nope,
                                 const float* __restrict__ pe, int rows, int C, float eps, int S, int skip,
                                 int pe_rows_per_t, int pe_T, int split, int reverse) {
  pdl_sync();  // launched with programmatic stream serialization (common.cuh)
  // each lane owns MAXV/2 groups of 4 consecutive channels: 16-byte fp32 loads, 8-byte bf16 stores
  constexpr int kQ = MAXV / 2;
  const int warps_per_block = blockDim.x >> 5;
  // reverse: last rows first.  The producer in front of a LayerNorm (a reduce-add GEMM walking its m-tiles upwards)
  // leaves the LAST rows it wrote in L2; starting there reads them before they are evicted.
  const int blk = reverse ? static_cast<int>(gridDim.x - 1 - blockIdx.x) : static_cast<int>(blockIdx.x);
  const int row = blk * warps_per_block + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int f = row / S, s = row - f * S;
  if (s < skip) return;
  const long long orow = static_cast<long long>(f) * (S - skip) + (s - skip);
  const TIn* xr = x + static_cast<long long>(row) * C;
  float4 v[kQ];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < kQ; ++i) {
    const int c = (i * 32 + lane) * 4;
    v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < C) {
      if constexpr (sizeof(TIn) == 4) {
        v[i] = *reinterpret_cast<const float4*>(xr + c);
      } else {
        const uint2 u = *reinterpret_cast<const uint2*>(xr + c);
        const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
        const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
        v[i] = make_float4(a.x, a.y, b.x, b.y);
      }
      sum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
  }
  const float mean = warp_sum(sum) / C;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < kQ; ++i) {
    const int c = (i * 32 + lane) * 4;
    if (c < C) {
      const float d0 = v[i].x - mean, d1 = v[i].y - mean, d2 = v[i].z - mean, d3 = v[i].w - mean;
      sq += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
    }
  }
  const float rstd = rsqrtf(warp_sum(sq) / C + eps);
  const float* per = pe ? pe + static_cast<long long>((row / pe_rows_per_t) % pe_T) * C : nullptr;
  const int ldo = split ? 3 * C : C;  // split outputs: [hi | lo | hi], parts C apart
  bf16* o = out + orow * ldo;
  bf16* o2 = out_nope ? out_nope + orow * ldo : nullptr;
#pragma unroll
  for (int i = 0; i < kQ; ++i) {
    const int c = (i * 32 + lane) * 4;
    if (c < C) {
      const float4 g = *reinterpret_cast<const float4*>(gamma + c);
      const float4 bb = *reinterpret_cast<const float4*>(beta + c);
      float r0 = (v[i].x - mean) * rstd * g.x + bb.x, r1 = (v[i].y - mean) * rstd * g.y + bb.y;
      float r2 = (v[i].z - mean) * rstd * g.z + bb.z, r3 = (v[i].w - mean) * rstd * g.w + bb.w;
      if (o2) {
        if (split) {
          st_pair(o2 + c, r0, r1, split), st_pair(o2 + c + 2, r2, r3, split);
        } else {
          const __nv_bfloat162 h0 = __floats2bfloat162_rn(r0, r1), h1 = __floats2bfloat162_rn(r2, r3);
          *reinterpret_cast<uint2*>(o2 + c) = make_uint2(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
        }
      }
      if (per) {
        const float4 pp = *reinterpret_cast<const float4*>(per + c);
        r0 += pp.x, r1 += pp.y, r2 += pp.z, r3 += pp.w;
      }
      if (split) {
        st_pair(o + c, r0, r1, split), st_pair(o + c + 2, r2, r3, split);
      } else {
        const __nv_bfloat162 h0 = __floats2bfloat162_rn(r0, r1), h1 = __floats2bfloat162_rn(r2, r3);
        *reinterpret_cast<uint2*>(o + c) = make_uint2(*reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
      }
    }
  }
}

// ------------------------------------------------------------------ GroupNorm (NHWC, 32 groups)
// stats: one block per (image, row chunk).  A thread owns 8 consecutive channels (one 16-byte load per row; the first
// version read one bf16 pair per thread and row and hit shared-memory atomics once per row) of every rpp-th row of the
// chunk, keeps its sums in registers, and at the end the block reduces them in a FIXED order to its 64 partials, which go
// into the fp64 sums of the image (64 global fp64 atomics per block, ~4 blocks per SM in the grid).
__global__ void __launch_bounds__(256) groupnorm_stats_kernel(const bf16* __restrict__ x, double* __restrict__ stats, int hw, int C,
                                                              int ldx, int rows_per_block, int split) {
  __shared__ float sh[64];
  __shared__ float part[256 * 8];        // per thread: (sum, sum of squares) of its 4 channel pairs
  const int img = blockIdx.y;
  const int r0 = blockIdx.x * rows_per_block;
  const int r1 = min(hw, r0 + rows_per_block);
  const int c8 = C >> 3;                 // threads per row
  const int rpp = 256 / c8;              // rows per pass of the block
  const int tr = threadIdx.x / c8, c = (threadIdx.x - tr * c8) << 3;
  const int cg = C / 32;
  if (tr < rpp) {
    float s[4] = {0.f, 0.f, 0.f, 0.f}, ss[4] = {0.f, 0.f, 0.f, 0.f};
    for (int r = r0 + tr; r < r1; r += rpp) {
      const bf16* src = x + (static_cast<long long>(img) * hw + r) * ldx + c;
      const uint4 hi = __ldg(reinterpret_cast<const uint4*>(src));
      uint4 lo = make_uint4(0, 0, 0, 0);
      if (split) lo = __ldg(reinterpret_cast<const uint4*>(src + split));
      const uint32_t* ph = &hi.x;
      const uint32_t* pl = &lo.x;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(ph + i));
        if (split) {
          const float2 l = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(pl + i));
          t.x += l.x, t.y += l.y;
        }
        s[i] += t.x + t.y;
        ss[i] += t.x * t.x + t.y * t.y;
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      part[(threadIdx.x * 4 + i) * 2] = s[i];
      part[(threadIdx.x * 4 + i) * 2 + 1] = ss[i];
    }
  }
  __syncthreads();
  // fixed-order reduction (shared-memory float atomics would make the sums, and through bf16 roundings the whole
  // frame, differ from run to run): thread (g, k) adds the pair partials of group g over rows and pairs in order
  if (threadIdx.x < 64) {
    const int g = threadIdx.x >> 1, k = threadIdx.x & 1;
    const int p0 = g * (cg >> 1), p1 = p0 + (cg >> 1);     // channel pairs of the group (cg is even)
    float acc = 0.f;
    for (int row = 0; row < rpp; ++row)
      for (int pr = p0; pr < p1; ++pr) acc += part[((row * c8 + (pr >> 2)) * 4 + (pr & 3)) * 2 + k];
    sh[threadIdx.x] = acc;
  }
  __syncthreads();
  if (threadIdx.x < 64) atomicAdd(&stats[static_cast<long long>(img) * 64 + threadIdx.x], static_cast<double>(sh[threadIdx.x]));
}

// mean / rstd per (image, group) from the fp64 sums, once (redoing the fp64 divisions in the apply pass was measured
// slower even at one pair of divisions per thread: gn74 81.8 -> 90.2 us).
__global__ void groupnorm_finalize_kernel(const double* __restrict__ stats, float2* __restrict__ mr, int total, int hw,
                                          int C, float eps) {
  pdl_sync();  // launched with programmatic stream serialization (common.cuh)
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const double n = static_cast<double>(hw) * (C / 32);
  const double mean = stats[2 * i] / n;
  const double var = stats[2 * i + 1] / n - mean * mean;
  mr[i] = make_float2(static_cast<float>(mean), rsqrtf(static_cast<float>(var > 0 ? var : 0) + eps));
}

// One thread per (pixel, 8 channels): 16-byte loads and stores; mean / rstd are looked up per channel pair unless
// C/32 % 8 == 0 (a pair never straddles a group, an 8-channel chunk may).  A block covers 256 / (C/8) consecutive pixels, so pixel and chunk come from two
// small 32-bit divisions of threadIdx (the first version divided a 64-bit linear index twice per thread).
__global__ void __launch_bounds__(256) groupnorm_apply_kernel(const bf16* __restrict__ x, const float2* __restrict__ mr,
                                                              const float* __restrict__ gamma, const float* __restrict__ beta,
                                                              bf16* __restrict__ out, int hw, int C, int ldx, long long total_oct,
                                                              int split) {
  pdl_sync();  // launched with programmatic stream serialization (common.cuh)
  const unsigned c8 = static_cast<unsigned>(C) >> 3;
  const unsigned ppb = 256u / c8;                      // pixels per block
  const unsigned tp = threadIdx.x / c8;
  if (tp >= ppb) return;
  const int c = static_cast<int>(threadIdx.x - tp * c8) << 3;
  const long long pix = static_cast<long long>(blockIdx.x) * ppb + tp;
  if (pix * c8 >= total_oct) return;
  const int img = static_cast<int>(static_cast<unsigned>(pix) / static_cast<unsigned>(hw));   // pixels < 2^32
  const int cg = C / 32;
  const bf16* src = x + pix * ldx + c;
  bf16* dst = out + pix * (split ? 3 * C : C) + c;
  const uint4 hi = __ldg(reinterpret_cast<const uint4*>(src));
  uint4 lo = make_uint4(0, 0, 0, 0);
  if (split) lo = __ldg(reinterpret_cast<const uint4*>(src + split));
  const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + c)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + c + 4));
  const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + c)), b1 = __ldg(reinterpret_cast<const float4*>(beta + c + 4));
  const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
  const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
  const uint32_t* ph = &hi.x;
  const uint32_t* pl = &lo.x;
  float r[8];
  float2 m = __ldg(mr + img * 32 + c / cg);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int cc = c + 2 * i;
    if (cg % 8 != 0 && i > 0) m = __ldg(mr + img * 32 + cc / cg);   // a chunk may straddle groups (cg = 2, 4, 12, 24 ...)
    float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(ph + i));
    if (split) {
      const float2 l = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(pl + i));
      t.x += l.x, t.y += l.y;
    }
    r[2 * i] = (t.x - m.x) * m.y * gg[2 * i] + bb[2 * i];
    r[2 * i + 1] = (t.y - m.x) * m.y * gg[2 * i + 1] + bb[2 * i + 1];
  }
  if (split) {
#pragma unroll
    for (int i = 0; i < 4; ++i) st_pair(dst + 2 * i, r[2 * i], r[2 * i + 1], C);  // split in -> split out, parts C apart
  } else {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(r[2 * i], r[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// ------------------------------------------------------------------ patch im2col (14x14, stride 14)
// x fp32 NCHW [frames,3,H,W] -> A bf16 [frames*gh*gw, ldk] with k = (c, kh, kw) (patch_embed.py:66).
__global__ void patch_im2col_kernel(const float* __restrict__ x, bf16* __restrict__ a, int H, int W, int gh, int gw,
                                    int ldk, int split) {
  pdl_sync();  // launched with programmatic stream serialization (common.cuh)
  const long long row = blockIdx.x;
  const int f = static_cast<int>(row / (gh * gw));
  const int p = static_cast<int>(row - static_cast<long long>(f) * gh * gw);
  const int ph = p / gw, pw = p - ph * gw;
  for (int k = threadIdx.x; k < ldk; k += blockDim.x) {
    float v = 0.f;
    if (k < 588) {
      const int c = k / 196, r = k - c * 196;
      const int kh = r / 14, kw = r - kh * 14;
      v = x[((static_cast<long long>(f) * 3 + c) * H + ph * 14 + kh) * W + pw * 14 + kw];
    }
    if (split) {  // row = [hi | lo | hi], parts ldk apart
      const bf16 hi = __float2bfloat16(v);
      bf16* o = a + row * 3 * ldk + k;
      o[0] = hi, o[ldk] = __float2bfloat16(v - __bfloat162float(hi)), o[2 * ldk] = hi;
    } else {
      a[row * ldk + k] = __float2bfloat16(v);
    }
  }
}

// tokens[f, 0] = cls + pos[0]; tokens[f, 1+i] = patch[f*hw+i] + pos[1+i]   (dinov2.py:218-219)
__global__ void add_cls_pos_kernel(const float* __restrict__ patch, const float* __restrict__ cls,
                                   const float* __restrict__ pos, float* __restrict__ tokens, int hw, int D,
                                   long long total4) {
  pdl_sync();  // launched with programmatic stream serialization (common.cuh)
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total4) return;
  const int d4 = D / 4;
  const long long row = idx / d4;
  const int c = static_cast<int>(idx - row * d4) * 4;
  const int S = hw + 1;
  const int f = static_cast<int>(row / S), s = static_cast<int>(row - static_cast<long long>(f) * S);
  float4 v = (s == 0) ? *reinterpret_cast<const float4*>(cls + c)
                      : *reinterpret_cast<const float4*>(patch + (static_cast<long long>(f) * hw + s - 1) * D + c);
  const float4 pp = *reinterpret_cast<const float4*>(pos + static_cast<long long>(s) * D + c);
  v.x += pp.x, v.y += pp.y, v.z += pp.z, v.w += pp.w;
  *reinterpret_cast<float4*>(tokens + row * D + c) = v;
}

// ------------------------------------------------------------------ bilinear, align_corners=True, NHWC bf16
// (util/blocks.py:156-158, dpt_temporal.py:116-118).  HBM-bound: algorithmic bytes = n*(h*w + H*W)*C*2.
// grid = (x-segments, y-strips, n); a block covers 256/(C/8) consecutive output columns, one thread per
// (column, 8 channels), and walks a strip of output rows keeping its two source rows in registers: a
// source row is fetched once per strip instead of once per output row (the one-output-per-thread
// version moved 4x the output bytes through L2 and ran at 2.4 TB/s).  32-bit index math only.
// Strip height is chosen per launch (32 / 16 / 8 rows: longer strips re-read fewer boundary rows; measured
// 148^2 -> 296^2 x 256: 359 / 335 / 328 us for 8 / 16 / 32), shorter when the grid would not fill the GPU.

__global__ void __launch_bounds__(256) bilinear_nhwc_kernel(const bf16* __restrict__ x, bf16* __restrict__ out, int h, int w,
                                                            int H, int W, int C, int ldx, int ldo, float sy, float sx, int strip) {
  pdl_sync();  // launched with programmatic stream serialization (common.cuh)
  const int c8 = C >> 3;
  const int px = threadIdx.x / c8;
  const int c = (threadIdx.x - px * c8) << 3;
  const int ppb = 256 / c8;
  const int X = blockIdx.x * ppb + px;
  if (X >= W || px >= ppb) return;
  const int n = blockIdx.z;
  const float fx = sx * X;
  const int x0 = min(static_cast<int>(fx), w - 1);
  const int x1 = min(x0 + 1, w - 1);
  const float lx = fx - x0;
  const bf16* base = x + static_cast<long long>(n) * h * w * ldx + c;
  const long long o0 = static_cast<long long>(x0) * ldx, o1 = static_cast<long long>(x1) * ldx;
  // Separable form: lx is fixed per thread, so each source row is reduced ONCE to its horizontal lerp
  // h(y) = x[y][x0] + lx * (x[y][x1] - x[y][x0]) (8 channels, fp32 registers) when it is loaded, and an
  // output row is one vertical lerp h0 + ly * (h1 - h0) of the two rows in registers.  The 4-tap form
  // unpacked all four bf16 taps again for every output row (~80 instructions per 16-byte store, which made
  // the kernel issue-bound at ~3.7 TB/s); this needs ~20 per store plus ~35 per source row.
  float h0[8], h1[8];
  auto load_row = [&](int y, float (&hh)[8]) {
    const bf16* r = base + static_cast<long long>(y) * w * ldx;
    const uint4 p = __ldg(reinterpret_cast<const uint4*>(r + o0));
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(r + o1));
    const uint32_t* pp = &p.x;
    const uint32_t* qq = &q.x;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 fp = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(pp + i));
      const float2 fq = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(qq + i));
      hh[2 * i] = fmaf(lx, fq.x - fp.x, fp.x);
      hh[2 * i + 1] = fmaf(lx, fq.y - fp.y, fp.y);
    }
  };
  int cur0 = -1, cur1 = -1;
  const int Yb = blockIdx.y * strip, Ye = min(Yb + strip, H);
  bf16* orow = out + ((static_cast<long long>(n) * H + Yb) * W + X) * ldo + c;
  for (int Y = Yb; Y < Ye; ++Y, orow += static_cast<long long>(W) * ldo) {
    const float fy = sy * Y;
    const int y0 = min(static_cast<int>(fy), h - 1);
    const int y1 = min(y0 + 1, h - 1);
    const float ly = fy - y0;
    if (y0 != cur0) {
      if (y0 == cur1) {
#pragma unroll
        for (int i = 0; i < 8; ++i) h0[i] = h1[i];
      } else {
        load_row(y0, h0);
      }
      cur0 = y0;
    }
    if (y1 != cur1) {
      if (y1 == y0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) h1[i] = h0[i];
      } else {
        load_row(y1, h1);
      }
      cur1 = y1;
    }
    uint32_t o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 r2 = __floats2bfloat162_rn(fmaf(ly, h1[2 * i] - h0[2 * i], h0[2 * i]),
                                                      fmaf(ly, h1[2 * i + 1] - h0[2 * i + 1], h0[2 * i + 1]));
      o[i] = *reinterpret_cast<const uint32_t*>(&r2);
    }
    __stcs(reinterpret_cast<uint4*>(orow), make_uint4(o[0], o[1], o[2], o[3]));
  }
}

// Split-tensor version (FP32-accuracy mode): taps are hi + lo, the interpolation runs in fp32 and
// the result is written back as [hi | lo | hi].  One thread per (output pixel, 8 channels).
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const uint32_t* w = &u.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(w + i));
    f[2 * i] = t.x, f[2 * i + 1] = t.y;
  }
}
__global__ void __launch_bounds__(256) bilinear_split_kernel(const bf16* __restrict__ x, bf16* __restrict__ out, int h, int w,
                                                             int H, int W, int C, int ldx, int ldo, int sin, int sout,
                                                             float sy, float sx) {
  const int c8 = C >> 3;
  const int px = threadIdx.x / c8;
  const int c = (threadIdx.x - px * c8) << 3;
  const int ppb = 256 / c8;
  const int X = blockIdx.x * ppb + px;
  if (X >= W || px >= ppb) return;
  const int Y = blockIdx.y, n = blockIdx.z;
  const float fy = sy * Y, fx = sx * X;
  const int y0 = min(static_cast<int>(fy), h - 1), x0 = min(static_cast<int>(fx), w - 1);
  const int y1 = min(y0 + 1, h - 1), x1 = min(x0 + 1, w - 1);
  const float ly = fy - y0, lx = fx - x0;
  const float wt[4] = {(1.f - ly) * (1.f - lx), (1.f - ly) * lx, ly * (1.f - lx), ly * lx};
  const int ys[4] = {y0, y0, y1, y1}, xs[4] = {x0, x1, x0, x1};
  const bf16* base = x + static_cast<long long>(n) * h * w * ldx + c;
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const bf16* src = base + (static_cast<long long>(ys[t]) * w + xs[t]) * ldx;
    float hi[8], lo[8];
    unpack8(__ldg(reinterpret_cast<const uint4*>(src)), hi);
    unpack8(__ldg(reinterpret_cast<const uint4*>(src + sin)), lo);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] += wt[t] * (hi[i] + lo[i]);
  }
  bf16* o = out + ((static_cast<long long>(n) * H + Y) * W + X) * ldo + c;
#pragma unroll
  for (int i = 0; i < 8; i += 2) st_pair(o + i, acc[i], acc[i + 1], sout);
}

// fp32 single-channel bilinear (video_depth.py:66,111-112), optional ReLU.
__global__ void bilinear_f32_kernel(const float* __restrict__ x, float* __restrict__ out, int h, int w, int H, int W,
                                    float sy, float sx, int relu, long long total) {
  pdl_sync();  // launched with programmatic stream serialization (common.cuh)
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int X = static_cast<int>(idx % W);
  const long long t = idx / W;
  const int Y = static_cast<int>(t % H);
  const long long n = t / H;
  const float fy = sy * Y, fx = sx * X;
  int y0 = min(static_cast<int>(fy), h - 1), x0 = min(static_cast<int>(fx), w - 1);
  const int y1 = min(y0 + 1, h - 1), x1 = min(x0 + 1, w - 1);
  const float ly = fy - y0, lx = fx - x0;
  const float* b = x + n * h * w;
  float v = (1.f - ly) * ((1.f - lx) * b[y0 * w + x0] + lx * b[y0 * w + x1]) +
            ly * ((1.f - lx) * b[y1 * w + x0] + lx * b[y1 * w + x1]);
  if (relu) v = fmaxf(v, 0.f);
  out[idx] = v;
}

// 3x3 stride-2 pad-1 im2col for dpt.py:85-90: x NHWC [n,h,w,C] -> A [n*ho*wo, 9*C], k = (kh,kw,c)
__global__ void im2col_s2_kernel(const bf16* __restrict__ x, bf16* __restrict__ a, int h, int w, int ho, int wo, int C,
                                 long long total) {
  pdl_sync();  // launched with programmatic stream serialization (common.cuh)
  const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int c8 = C / 8;
  const int c = static_cast<int>(idx % c8) * 8;
  long long t = idx / c8;
  const int tap = static_cast<int>(t % 9);
  t /= 9;
  const int X = static_cast<int>(t % wo);
  t /= wo;
  const int Y = static_cast<int>(t % ho);
  const int n = static_cast<int>(t / ho);
  const int kh = tap / 3, kw = tap - kh * 3;
  const int iy = Y * 2 - 1 + kh, ix = X * 2 - 1 + kw;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (iy >= 0 && iy < h && ix >= 0 && ix < w)
    v = __ldg(reinterpret_cast<const uint4*>(x + ((static_cast<long long>(n) * h + iy) * w + ix) * C + c));
  const long long orow = (static_cast<long long>(n) * ho + Y) * wo + X;
  *reinterpret_cast<uint4*>(a + orow * (9LL * C) + tap * C + c) = v;
}

}  // namespace vda

using namespace vda;

template <typename TIn>
static int launch_ln(const void* x, const float* gamma, const float* beta, void* out, void* out_nope, const float* pe,
                     int rows, int C, float eps, int S, int skip, int pe_rows_per_t, int pe_T, int split, cudaStream_t st) {
  const int warps = 8;
  dim3 grid((rows + warps - 1) / warps), block(warps * 32);
  static const int reverse = [] { const char* e = getenv("VDA_LN_REVERSE"); return e ? atoi(e) : 1; }();  // A/B knob (profiles/r02f_ab_VDA_LN_REVERSE.log)
  cudaError_t le = cudaSuccess;
#define VDA_LN(MAXV)                                                                                              \
  le = launch_pdl(layernorm_kernel<TIn, MAXV>, grid, block, 0, st, static_cast<const TIn*>(x), gamma, beta,       \
                  static_cast<bf16*>(out), static_cast<bf16*>(out_nope), pe, rows, C, eps, S, skip, pe_rows_per_t, \
                  pe_T, split ? C : 0, reverse)
  if (C <= 256) VDA_LN(4);
  else if (C <= 512) VDA_LN(8);
  else if (C <= 1024) VDA_LN(16);
  else return set_error(-1, "layernorm: C=%d > 1024 unsupported", C);
#undef VDA_LN
  return check_cuda(le, "layernorm launch");
}

// ------------------------------------------------------------------ allocation-free GroupNorm scratch
// Scratch of vda_groupnorm_nhwc_static: lives in the module image of every device that loads the library (48 KB
// each, until exit).  In HOST code the name `g_gn_static_ws` is only a handle for the runtime, not a device address:
// the address is asked for once per device, with that device current, through cudaGetSymbolAddress.
constexpr int kGnStaticMaxN = VDA_GROUPNORM_STATIC_MAX_N;
static __device__ __align__(256) double g_gn_static_ws[96 * kGnStaticMaxN];

static int gn_static_ws(double** out) {
  static double* resolved[kMaxDevices] = {};  // benign race: every thread resolves the same address
  const int dev = current_device();
  if (dev < 0 || dev >= kMaxDevices) return set_error(-2, "vda_groupnorm_nhwc_static: device %d out of range", dev);
  if (!resolved[dev]) {
    double* p = nullptr;
    const int rc = check_cuda(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_gn_static_ws),
                              "cudaGetSymbolAddress(g_gn_static_ws)");
    if (rc) return rc;
    if (!p) return set_error(-2, "vda_groupnorm_nhwc_static: no device address for the static scratch on device %d", dev);
    resolved[dev] = p;
  }
  *out = resolved[dev];
  return 0;
}

extern "C" {

int vda_layernorm(const void* x, int x_is_f32, const float* gamma, const float* beta, void* out, void* out_nope,
                  const float* pe, int rows, int C, float eps, int tokens_per_frame, int skip, int pe_rows_per_t,
                  int pe_T, int split, void* stream) {
  if (!x || !gamma || !beta || !out || rows <= 0 || C % 4) return set_error(-1, "vda_layernorm: bad arguments (C%%4)");
  if (tokens_per_frame <= 0) tokens_per_frame = rows, skip = 0;
  if (pe && (pe_rows_per_t <= 0 || pe_T <= 0)) return set_error(-1, "vda_layernorm: pe needs rows_per_t and T");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return x_is_f32 ? launch_ln<float>(x, gamma, beta, out, out_nope, pe, rows, C, eps, tokens_per_frame, skip,
                                     pe_rows_per_t, pe_T, split, st)
                  : launch_ln<bf16>(x, gamma, beta, out, out_nope, pe, rows, C, eps, tokens_per_frame, skip,
                                    pe_rows_per_t, pe_T, split, st);
}

int vda_groupnorm_nhwc(const void* x, int ldx, const float* gamma, const float* beta, void* out, double* stats_ws,
                       int n, int hw, int C, float eps, int split, void* stream) {
  if (!x || !out || !stats_ws || C % 64 || ldx % 8 || split % 8 || n <= 0 || hw <= 0)
    return set_error(-1, "vda_groupnorm_nhwc: bad arguments (C%%64, ldx%%8)");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = check_cuda(cudaMemsetAsync(stats_ws, 0, sizeof(double) * 64 * n, st), "groupnorm memset");
  if (rc) return rc;
  if (C > 2048) return set_error(-1, "vda_groupnorm_nhwc: C <= 2048");
  // ~4 blocks per SM over the whole launch, at least one pass of rows per block
  const int rpp = 256 / (C / 8);
  int chunks = (4 * num_sms() + n - 1) / n;
  if (chunks > (hw + rpp - 1) / rpp) chunks = (hw + rpp - 1) / rpp;
  if (chunks < 1) chunks = 1;
  const int rows_per_block = (hw + chunks - 1) / chunks;
  dim3 grid((hw + rows_per_block - 1) / rows_per_block, n);
  groupnorm_stats_kernel<<<grid, 256, 0, st>>>(static_cast<const bf16*>(x), stats_ws, hw, C, ldx, rows_per_block, split);
  // stats_ws: 64*n doubles of sums, the 32*n float2 (mean, rstd) pairs behind them
  float2* mr = reinterpret_cast<float2*>(stats_ws + 64 * n);
  rc = check_cuda(cudaGetLastError(), "groupnorm stats launch");  // plain launch: it follows a memset node
  if (rc) return rc;
  rc = check_cuda(launch_pdl(groupnorm_finalize_kernel, dim3((32 * n + 127) / 128), dim3(128), 0, st, stats_ws, mr, 32 * n, hw,
                             C, eps),
                  "groupnorm finalize launch");
  if (rc) return rc;
  const long long total = static_cast<long long>(n) * hw * (C / 8);
  if (static_cast<long long>(n) * hw >= (1LL << 32)) return set_error(-1, "vda_groupnorm_nhwc: n * hw < 2^32");
  const long long ppb = 256 / (C / 8);
  return check_cuda(launch_pdl(groupnorm_apply_kernel, dim3(static_cast<unsigned>((static_cast<long long>(n) * hw + ppb - 1) / ppb)), dim3(256), 0, st,
                               static_cast<const bf16*>(x), mr, gamma, beta, static_cast<bf16*>(out), hw, C, ldx, total, split),
                    "groupnorm launch");
}

int vda_patch_im2col(const float* x, void* a, int frames, int H, int W, int ldk, int split, void* stream) {
  if (!x || !a || H % 14 || W % 14 || ldk < 588 || ldk % 8) return set_error(-1, "vda_patch_im2col: bad arguments");
  const int gh = H / 14, gw = W / 14;
  return check_cuda(launch_pdl(patch_im2col_kernel, dim3(frames * gh * gw), dim3(160), 0, static_cast<cudaStream_t>(stream), x,
                               static_cast<bf16*>(a), H, W, gh, gw, ldk, split),
                    "patch_im2col launch");
}

int vda_add_cls_pos(const float* patch, const float* cls, const float* pos, float* tokens, int frames, int hw, int D,
                    void* stream) {
  if (!patch || !cls || !pos || !tokens || D % 4) return set_error(-1, "vda_add_cls_pos: bad arguments");
  const long long total = static_cast<long long>(frames) * (hw + 1) * (D / 4);
  return check_cuda(launch_pdl(add_cls_pos_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), 0,
                               static_cast<cudaStream_t>(stream), patch, cls, pos, tokens, hw, D, total),
                    "add_cls_pos launch");
}

int vda_bilinear_nhwc(const void* x, int ldx, void* out, int ldo, int n, int h, int w, int H, int W, int C,
                      int split_in, int split_out, void* stream) {
  if (!x || !out || C % 8 || ldx % 8 || ldo % 8) return set_error(-1, "vda_bilinear_nhwc: bad arguments");
  if ((split_in == 0) != (split_out == 0) || split_in % 8 || split_out % 8)
    return set_error(-1, "vda_bilinear_nhwc: split_in / split_out must both be set (multiples of 8) or both 0");
  const float sy = H > 1 ? static_cast<float>(h - 1) / (H - 1) : 0.f;
  const float sx = W > 1 ? static_cast<float>(w - 1) / (W - 1) : 0.f;
  if (C > 2048 || H > 65535 || n > 65535) return set_error(-1, "vda_bilinear_nhwc: C <= 2048, H, n <= 65535");
  const int ppb = 256 / (C / 8);
  if (split_in) {
    const dim3 grid((W + ppb - 1) / ppb, H, n);
    bilinear_split_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const bf16*>(x), static_cast<bf16*>(out), h, w, H, W, C, ldx, ldo, split_in, split_out, sy, sx);
    return check_cuda(cudaGetLastError(), "bilinear_split launch");
  }
  int strip = 32;
  while (strip > 8 && static_cast<long long>((W + ppb - 1) / ppb) * ((H + strip - 1) / strip) * n < 2048) strip >>= 1;
  const dim3 grid((W + ppb - 1) / ppb, (H + strip - 1) / strip, n);
  return check_cuda(launch_pdl(bilinear_nhwc_kernel, grid, dim3(256), 0, static_cast<cudaStream_t>(stream),
                               static_cast<const bf16*>(x), static_cast<bf16*>(out), h, w, H, W, C, ldx, ldo, sy, sx, strip),
                    "bilinear_nhwc launch");
}

int vda_bilinear_f32(const float* x, float* out, int n, int h, int w, int H, int W, int relu, void* stream) {
  if (!x || !out) return set_error(-1, "vda_bilinear_f32: bad arguments");
  const float sy = H > 1 ? static_cast<float>(h - 1) / (H - 1) : 0.f;
  const float sx = W > 1 ? static_cast<float>(w - 1) / (W - 1) : 0.f;
  const long long total = static_cast<long long>(n) * H * W;
  return check_cuda(launch_pdl(bilinear_f32_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), 0,
                               static_cast<cudaStream_t>(stream), x, out, h, w, H, W, sy, sx, relu, total),
                    "bilinear_f32 launch");
}

int vda_im2col_s2(const void* x, void* a, int n, int h, int w, int C, void* stream) {
  if (!x || !a || C % 8) return set_error(-1, "vda_im2col_s2: bad arguments");
  const int ho = (h - 1) / 2 + 1, wo = (w - 1) / 2 + 1;
  const long long total = static_cast<long long>(n) * ho * wo * 9 * (C / 8);
  return check_cuda(launch_pdl(im2col_s2_kernel, dim3(static_cast<unsigned>((total + 255) / 256)), dim3(256), 0,
                               static_cast<cudaStream_t>(stream), static_cast<const bf16*>(x), static_cast<bf16*>(a), h, w, ho, wo, C,
                               total),
                    "im2col_s2 launch");
}

// GroupNorm without any allocation by the caller or the library: the statistics scratch is the file-scope device array
// above.  Same launches as vda_groupnorm_nhwc (the sums are cleared by its memset each call; the array otherwise keeps
// the previous call's contents).  All calls on one device share the array: calls that overlap on the GPU race.
int vda_groupnorm_nhwc_static(const void* x, int ldx, const float* gamma, const float* beta, void* out, int n, int hw,
                              int C, float eps, int split, void* stream) {
  if (n > kGnStaticMaxN)
    return set_error(-8, "vda_groupnorm_nhwc_static: n=%d images exceed the static scratch (VDA_GROUPNORM_STATIC_MAX_N=%d); "
                         "use vda_groupnorm_nhwc with caller-owned scratch", n, kGnStaticMaxN);
  if (!x || !out || C % 64 || n <= 0 || hw <= 0) return set_error(-1, "vda_groupnorm_nhwc_static: bad arguments (C%%64)");
  double* ws = nullptr;
  const int rc = gn_static_ws(&ws);
  if (rc) return rc;
  return vda_groupnorm_nhwc(x, ldx, gamma, beta, out, ws, n, hw, C, eps, split, stream);
}

}  // extern "C"

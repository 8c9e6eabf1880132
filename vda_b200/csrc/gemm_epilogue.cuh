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
#ifndef VDA_EPI_WARPS
#define VDA_EPI_WARPS 8  // measured: 16 warps (96 regs) is not faster than 8 (126 regs)
#endif
constexpr int kEpiWarps = VDA_EPI_WARPS;  // 4 TMEM lane quadrants x kEpiParts column slices
static_assert(VDA_EPI_WARPS == 8, "the box epilogues (kRedBoxes boxes x 2 column slices, named barriers 2/3) are laid out for 8 epilogue warps");
constexpr int kEpiParts = kEpiWarps / 4;
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

// FP32-accuracy mode ("split" activations): an fp32 value x is carried as two bf16 numbers
// hi = bf16(x), lo = bf16(x - hi) (|x - hi - lo| <= 2^-17 |x|), laid out [hi | lo | hi] with
// `stride` elements between the parts so that the tensor is directly the A operand of a GEMM whose
// weights are packed [Wh | Wh | Wl] along K:  A.W = hi.Wh + lo.Wh + hi.Wl  (the lo.Wl term is 2^-16).
__device__ __forceinline__ void store_split16(bf16* dst, const float (&v)[16], int stride) {
  float lo[16];
  uint32_t w[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    w[i] = *reinterpret_cast<uint32_t*>(&h);
    const float2 f = __bfloat1622float2(h);
    lo[2 * i] = v[2 * i] - f.x, lo[2 * i + 1] = v[2 * i + 1] - f.y;
  }
  uint4* d = reinterpret_cast<uint4*>(dst);
  d[0] = make_uint4(w[0], w[1], w[2], w[3]);
  d[1] = make_uint4(w[4], w[5], w[6], w[7]);
  d = reinterpret_cast<uint4*>(dst + 2 * stride);
  d[0] = make_uint4(w[0], w[1], w[2], w[3]);
  d[1] = make_uint4(w[4], w[5], w[6], w[7]);
  store_bf16x16(dst + stride, lo);
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
__device__ __forceinline__ void stage_bias_scale(const GemmParams& p, int n_t, int et, float* s_bias, float* s_scale,
                                                 bool with_bias = true) {
  if (et < BN) {
    const int col = n_t * BN + et;
    const bool in = col < p.N;
    s_bias[et] = (p.bias && in && with_bias) ? __ldg(p.bias + col) : 0.f;
    s_scale[et] = (p.scale && in) ? __ldg(p.scale + col) : 1.f;
  }
  named_bar_sync(1, kEpiThreads);
}

// One 16-column group of one row: v already holds acc + bias.
__device__ __forceinline__ void finish_group16(const GemmParams& p, const RowMap& rm, int col, float (&v)[16],
                                               const float* s_scale16, const float4* res_pre) {
  if (p.act == ACT_GELU) {
#pragma unroll
    for (int i = 0; i < 16; i += 2) gelu_fast2(v[i], v[i + 1]);
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
  // fp32 operands may have their own row pitch (split mode: the bf16 tensors are three parts wide)
  const long long off32 = (p.ld32 && p.store_mode == ST_PLAIN) ? rm.row * p.ld32 + col : off;
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
    load_f32x16(p.res_f32 + off32, t);
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += t[i];
  }
  if (p.res1) add_bf16x16(p.res1 + off, v);
  if (p.res2) add_bf16x16(p.res2 + off, v);
  if (p.split) {  // residuals are split tensors too: add their lo parts
    if (p.res1) add_bf16x16(p.res1 + off + p.split, v);
    if (p.res2) add_bf16x16(p.res2 + off + p.split, v);
  }
  if (p.out_f32) {
    float4* d = reinterpret_cast<float4*>(p.out_f32 + off32);
#pragma unroll
    for (int i = 0; i < 4; ++i) d[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  }
  if (p.split) {
    if (p.out) store_split16(p.out + off, v, p.split);
    if (p.out_relu) {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
      store_split16(p.out_relu + off, v, p.split);
    }
    return;
  }
  if (p.out) store_bf16x16(p.out + off, v);
  if (p.out_relu) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
    store_bf16x16(p.out_relu + off, v);
  }
}

// Columns [cbeg, cbeg + kPer) of the tile belong to column slice `part` (0..kEpiParts-1).
template <int BN>
struct EpiSlice {
  static constexpr int kPer = (BN / kEpiParts >= 32) ? BN / kEpiParts : 32;
};

// Requests the fp32 residual lines this thread will read for the tile into L2 (no registers held).
template <int BN>
__device__ __forceinline__ void epilogue_prefetch(const GemmParams& p, int m_tile, int n_t, int r, int half) {
  constexpr int kPer = EpiSlice<BN>::kPer;
  const int cbeg = half * kPer;
  if (cbeg >= BN) return;
  if (p.epi_tma == 5) {   // bf16 residual rows of the tile (cooperative boxes): 2 * kPer bytes per row and tensor
    if (p.res1 == nullptr && p.res2 == nullptr) return;
    const RowMap rm = map_row(p, m_tile, r);
    const int col = n_t * BN + cbeg;
    if (!rm.valid || col >= p.N) return;
#pragma unroll
    for (int i = 0; i < kPer; i += 64) {
      if (p.res1) prefetch_l2(p.res1 + rm.row * p.ldo + col + i);
      if (p.res2) prefetch_l2(p.res2 + rm.row * p.ldo + col + i);
    }
    return;
  }
  if (p.res_f32 == nullptr || p.store_mode != ST_PLAIN) return;
  const RowMap rm = map_row(p, m_tile, r);
  if (!rm.valid) return;
  const int col = n_t * BN + cbeg;
  const float* src = p.res_f32 + rm.row * (p.ld32 ? p.ld32 : p.ldo) + col;
#pragma unroll
  for (int i = 0; i < kPer; i += 32)
    if (col + i < p.N) prefetch_l2(src + i);
}

// ---------------------------------------------------------------------------------------------
// fp32 residual staging (proj / fc2 / motion-module epilogues: x += gamma * (acc + b), in place).
// Reading the residual with plain loads put a full L2/HBM round trip on every 16-column group of
// every thread (proj: 184 us against 82 us without the read).  Instead each epilogue thread streams
// its residual segments through a private ring of shared-memory slots with cp.async, kRing-1 groups
// (and across tile boundaries) ahead of their use; slots are thread-private, so no barrier is needed.
// The ring lives in the smem of GEMM stages that this launch leaves unused (GemmParams::stages).
// A work unit of the persistent loop: a whole tile, or one K range of a tail tile (GemmParams::ksplit).
struct WorkUnit {
  int tile, kb0, kb1;
};
__device__ __forceinline__ WorkUnit decode_unit(const GemmParams& p, int v) {
  WorkUnit u;
  if (p.ksplit <= 1 || v < p.tail_begin) {
    u.tile = v, u.kb0 = 0, u.kb1 = p.num_k_blocks;
  } else {
    const int t = v - p.tail_begin;
    const int q = t / p.ksplit, part = t - q * p.ksplit;
    u.tile = p.tail_begin + q;
    u.kb0 = part * p.num_k_blocks / p.ksplit;
    u.kb1 = (part + 1) * p.num_k_blocks / p.ksplit;
  }
  return u;
}

struct TileSched {
  int begin, step, end;  // this CTA's tile sequence
  int m_mul, m_add;      // m_tile = m_mul * (tile / n_tiles) + m_add  (2, rank for CTA pairs)
};

struct ResRing {
  uint32_t base;  // shared-space address of slot 0 of this thread (slot pitch = kEpiThreads * 64 B)
  int nbuf;       // slots in the ring; 0 = staging off
  int issued;     // groups issued so far (global group counter of this thread)
};
constexpr uint32_t kResSlotPitch = kEpiThreads * 64;

template <int BN>
__device__ __forceinline__ void res_ring_issue(const GemmParams& p, const TileSched& ts, ResRing& rr, int r, int half) {
  constexpr int kPer = EpiSlice<BN>::kPer;
  constexpr int kGroups = kPer / 16;
  const int k = rr.issued++;
  const int ti = k / kGroups, g = k - ti * kGroups;
  const int tile = ts.begin + ti * ts.step;
  if (tile < ts.end) {
    const int mq = tile / p.n_tiles, n_t = tile - mq * p.n_tiles;
    const RowMap rm = map_row(p, ts.m_mul * mq + ts.m_add, r);
    const int col = n_t * BN + half * kPer + 16 * g;
    if (rm.valid && col < p.N) {
      const float* src = p.res_f32 + rm.row * (p.ld32 ? p.ld32 : p.ldo) + col;
      const uint32_t dst = rr.base + static_cast<uint32_t>(k % rr.nbuf) * kResSlotPitch;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16 * i), "l"(src + 4 * i) : "memory");
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}

// Builds the ring in `bytes` of free smem at `area` and issues the first nbuf-1 groups.
template <int BN>
__device__ __forceinline__ ResRing make_res_ring(const GemmParams& p, const TileSched& ts, uint8_t* area, int bytes, int et,
                                                 int r, int half) {
  ResRing rr;
  rr.base = smem_u32(area) + static_cast<uint32_t>(et) * 64;
  rr.nbuf = 0;
  rr.issued = 0;
  constexpr int kPer = EpiSlice<BN>::kPer;
  if (!p.epi_tma && p.res_f32 != nullptr && p.store_mode == ST_PLAIN && p.act != ACT_GEGLU && half * kPer < BN) {
    const int n = bytes / static_cast<int>(kResSlotPitch);
    rr.nbuf = n >= 4 ? 4 : (n >= 3 ? 3 : (n >= 2 ? 2 : 0));
    for (int i = 0; i + 1 < rr.nbuf; ++i) res_ring_issue<BN>(p, ts, rr, r, half);
  }
  return rr;
}

// ---------------------------------------------------------------------------------------------
// In-place fp32 residual update  x += scale * act(acc + bias)  as a TMA reduce-add.
// With one thread per accumulator row, every warp-level global access of the epilogue touches 32
// different 128-byte lines, and L1TEX serialises those (about 16 B/clk/SM): reading and writing the
// fp32 residual cost more than the MMAs (ncu / A-B runs: proj 184 us vs 82 us without them).  Here
// the threads only write their delta rows into a swizzled shared-memory box (16 fp32 columns x 128
// rows); one elected thread per column slice hands the box to the TMA unit, which adds it onto the
// residual in L2 (cp.reduce.async.bulk.tensor .add).  The SM never reads the residual at all.
// epi_boxes(MODE) boxes per column slice rotate (gemm_params.h); one named barrier per box (ordering argument inline).
constexpr int kRedBoxBytes = 16 * 4 * kBlockM;  // 8 KB

__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, uint32_t src_smem, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(src_smem), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, uint32_t src_smem, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(src_smem), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, uint32_t src_smem, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(src_smem), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}

// Output element offset of every row of a tile for the scattered stores of MODE 3 (-1 = row not stored),
// computed once per tile by the first 128 epilogue threads into a table double-buffered by tile parity.
__device__ __forceinline__ void stage_row_offsets(const GemmParams& p, int m_tile, int et, long long* s_rowoff) {
  if ((p.epi_tma != 3 && p.epi_tma != 5) || et >= kBlockM) return;
  const RowMap rm = map_row(p, m_tile, et);
  long long off = -1;
  if (rm.valid && p.epi_tma == 5) {
    off = rm.row * p.ldo;   // plain addressing (GEMM rows or conv pixels)
  } else if (rm.valid) {
    // ST_CONVT: input pixel (n, ih, iw) -> output pixel (n, ih*s, iw*s); the tap (kh, kw) is added per box
    const int hw = p.ct_h * p.ct_w;
    const int m = static_cast<int>(rm.row);
    const int n = m / hw, rem = m - n * hw;
    const int ih = rem / p.ct_w, iw = rem - ih * p.ct_w;
    off = ((static_cast<long long>(n) * p.ct_h * p.ct_s + ih * p.ct_s) * (p.ct_w * p.ct_s) + iw * p.ct_s) * p.ldo;
  }
  s_rowoff[et] = off;
}

// Box epilogues.  MODE 1: fp32 reduce-add, 16 columns per box.  MODE 2: bf16 store, 32 columns per
// box (same 64-byte rows, SWIZZLE_64B; conv tiles use a 4-D map whose box is the tile's pixel patch).
// MODE 3: bf16 box as MODE 2, but written out by the slice's 128 threads themselves, four lanes per
// 64-byte row segment, to rows given by a per-tile offset table: for stores TMA cannot express as one box
// (ConvTranspose pixel shuffle).  A warp store then touches 8 lines instead of 32 (thread = row).
// `box_base`: shared address of this column slice's boxes; `box_ctr`: running box counter of the slice.
template <int BN, int MODE>
__device__ __forceinline__ void epilogue_tile_box(const GemmParams& p, int m_tile, int n_t, uint32_t t_acc, int r, int half,
                                                  const float* s_bias, const float* s_scale, uint32_t box_base, int& box_ctr,
                                                  bool issuer, const long long* s_rowoff = nullptr) {
  constexpr int kPer = EpiSlice<BN>::kPer;
  constexpr int kRedBoxes = epi_boxes(MODE);
  constexpr int kCols = MODE == 1 ? 16 : (MODE == 4 ? 64 : 32);  // accumulator columns per box
  const int cbeg = half * kPer;
  if (cbeg >= BN) return;
  // The accumulator chunk of box k+1 is requested as soon as the registers of chunk k have been packed, so
  // its TMEM latency runs under the shared-memory writes, the barrier and the TMA issue of box k (with a
  // K = 256 GEMM the MMAs of a tile take 2048 clk and eight strictly serial boxes did not fit under them).
  uint32_t raw[MODE == 4 ? 32 : kCols];
  __syncwarp();
  if constexpr (MODE == 1) {
    tmem_ld16(t_acc + cbeg, raw);
  } else {
    tmem_ld32(t_acc + cbeg, raw);
  }
#pragma unroll 1
  for (int c0 = cbeg; c0 < cbeg + kPer; c0 += kCols) {
    const int col = n_t * BN + c0;
    tmem_ld_wait();
    uint32_t pk[16];  // 64 bytes of this row: 16 fp32 or 32 bf16
    if constexpr (MODE == 4) {
      // GEGLU: accumulator columns [32j, 32j+16) = value, [32j+16, 32j+32) = gate -> 16 outputs; a box
      // takes 64 accumulator columns = 32 bf16 outputs, at output column col / 2 (weight rows are packed
      // in 16-row value / gate groups, vda_b200/packing.py)
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        float b[16], bg[16];
        lds_f32x16(s_bias + c0 + 32 * g, b);
        lds_f32x16(s_bias + c0 + 32 * g + 16, bg);
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
          float g0 = __uint_as_float(raw[16 + i]) + bg[i], g1 = __uint_as_float(raw[16 + i + 1]) + bg[i + 1];
          gelu_fast2(g0, g1);
          __nv_bfloat162 h2 = __floats2bfloat162_rn((__uint_as_float(raw[i]) + b[i]) * g0,
                                                    (__uint_as_float(raw[i + 1]) + b[i + 1]) * g1);
          pk[8 * g + (i >> 1)] = *reinterpret_cast<uint32_t*>(&h2);
        }
        if (g == 0 || c0 + kCols < cbeg + kPer) {  // second half of this box / first half of the next
          __syncwarp();
          tmem_ld32(t_acc + c0 + 32 * (g + 1), raw);
          if (g == 0) tmem_ld_wait();
        }
      }
    }
#pragma unroll
    for (int g = 0; g < (MODE == 4 ? 0 : kCols / 16); ++g) {
      float b[16], v[16];
      lds_f32x16(s_bias + c0 + 16 * g, b);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(raw[16 * g + i]) + b[i];
      if (p.act == ACT_GELU) {
#pragma unroll
        for (int i = 0; i < 16; i += 2) gelu_fast2(v[i], v[i + 1]);
      } else if (p.act == ACT_RELU) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
      }
      if (p.scale) {
        float sc[16];
        lds_f32x16(s_scale + c0 + 16 * g, sc);
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] *= sc[i];
      }
      if constexpr (MODE == 1) {
#pragma unroll
        for (int i = 0; i < 16; ++i) pk[i] = __float_as_uint(v[i]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          __nv_bfloat162 h2 = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
          pk[8 * g + i] = *reinterpret_cast<uint32_t*>(&h2);
        }
      }
    }
    if (MODE != 4 && c0 + kCols < cbeg + kPer) {  // raw is dead: fetch the next chunk behind the rest of this box
      __syncwarp();
      if constexpr (MODE == 1) {
        tmem_ld16(t_acc + c0 + kCols, raw);
      } else if constexpr (MODE != 4) {
        tmem_ld32(t_acc + c0 + kCols, raw);
      }
    }
    // box (box_ctr % kRedBoxes) is free: the store that last used it was confirmed read by the issuer before the
    // issuer joined the barrier of the previous box, which every thread has passed.
    const uint32_t box = box_base + static_cast<uint32_t>(box_ctr % kRedBoxes) * kRedBoxBytes;
    const uint32_t rowaddr = box + static_cast<uint32_t>(r) * 64;
    const uint32_t sw = (static_cast<uint32_t>(r) >> 1) & 3;
#pragma unroll
    for (int c = 0; c < 4; ++c)
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(rowaddr + ((static_cast<uint32_t>(c) ^ sw) << 4)),
                   "r"(pk[4 * c]), "r"(pk[4 * c + 1]), "r"(pk[4 * c + 2]), "r"(pk[4 * c + 3])
                   : "memory");
    if constexpr (MODE == 3) {
      named_bar_sync(2 + half, 128);
      if (col < p.N) {
        const int gg = col / p.ct_cout, co = col - gg * p.ct_cout;
        const int kh = gg / p.ct_s, kw = gg - kh * p.ct_s;
        const long long tapoff = (static_cast<long long>(kh) * (p.ct_w * p.ct_s) + kw) * p.ldo + co;
        const uint32_t chunk = static_cast<uint32_t>(r) & 3;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t row = (static_cast<uint32_t>(r) >> 2) + 32 * i;
          const long long ro = s_rowoff[row];
          uint4 v;
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                       : "r"(box + row * 64 + ((chunk ^ ((row >> 1) & 3)) << 4)));
          if (ro >= 0) *reinterpret_cast<uint4*>(p.out + ro + tapoff + chunk * 8) = v;
        }
      }
      // box reuse: a thread writes box k+2 only after the barrier of box k+1, which every thread joins
      // after it has finished copying box k out
      ++box_ctr;
      continue;
    }
    fence_proxy_async_smem();
    if (kRedBoxes == 2 && issuer) bulk_wait_read_all();  // box k-1 has left shared memory: whoever passes the barrier may reuse it
    named_bar_sync(2 + half, 128);
    if (issuer) {
      if (col < p.N) {  // rows >= M / cols >= N / pixels outside the map are clipped by TMA
        if constexpr (MODE == 1) {
          if (p.f32_store)
            tma_store_2d(&p.tmD, box, col, m_tile * kBlockM);  // plain fp32 output, no residual
          else
            tma_reduce_add_2d(&p.tmD, box, col, m_tile * kBlockM);
        } else if (p.a_mode == A_CONV3X3) {
          const int tpi = p.tiles_w * p.tiles_h;
          const int img = m_tile / tpi;
          const int t = m_tile - img * tpi;
          const int ty = t / p.tiles_w, tx = t - ty * p.tiles_w;
          tma_store_4d(&p.tmD, box, col, tx << p.tw_log2, ty * (kBlockM >> p.tw_log2), img);
        } else {
          tma_store_2d(&p.tmD, box, MODE == 4 ? (col >> 1) : col, m_tile * kBlockM);
        }
      }
      bulk_commit();
      if (kRedBoxes != 2) bulk_wait_read_1();  // 3 boxes: the previous box of this slice has been read out of shared memory
    }
    ++box_ctr;
  }
}

// ---------------------------------------------------------------------------------------------
// MODE 5, "cooperative boxes": bf16 output with bf16 residuals and / or a ReLU'd second output
// (out = acc + bias (+ res1 + res2), out_relu = relu(out): the ResidualConvUnit / FeatureFusion epilogues, util/blocks.py:78-91,
// 146-160, and the motion module's proj_out + residual).  The element-wise path reads and writes those tensors with
// thread = row, 32 different lines per warp access (conv 148x148 with two residuals: 887 us against 570-600 us without).
// Here every tensor moves through 64-byte-row boxes in shared memory like MODE 3: the slice's 128 threads fetch the
// residual boxes with cp.async, four lanes per 64-byte row run (8 lines per warp access), ONE BOX AHEAD (the first box of
// a tile is requested by coop_tile_begin before the wait for the accumulator); a thread then adds its own row, writes the
// result over the res1 box (ReLU'd copy over the res2 box) and the slice copies both out the same cooperative way.
// Per slice: 2 slots x {res1 | out, res2 | out_relu} boxes of 8 KB.  Barrier A of box k (residuals landed) is also the
// point where every thread has finished copying out box k-1, so slot k+1 (= slot k-1) may be refilled behind it.
__device__ __forceinline__ void coop_res_issue(const GemmParams& p, const long long* s_rowoff, int col, uint32_t box_a,
                                               uint32_t box_b, int r) {
  if ((p.res1 || p.res2) && col < p.N) {
    const uint32_t chunk = static_cast<uint32_t>(r) & 3;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t row = (static_cast<uint32_t>(r) >> 2) + 32 * i;
      const long long ro = s_rowoff[row];
      if (ro >= 0) {
        const uint32_t dst = row * 64 + ((chunk ^ ((row >> 1) & 3)) << 4);
        const long long src = ro + col + chunk * 8;
        if (p.res1) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(box_a + dst), "l"(p.res1 + src) : "memory");
        if (p.res2) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(box_b + dst), "l"(p.res2 + src) : "memory");
      }
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}

// Called by every epilogue thread at the top of a tile (row-offset table of the tile staged and visible).
template <int BN>
__device__ __forceinline__ void coop_tile_begin(const GemmParams& p, int n_t, int r, int half, uint32_t box_base, int box_ctr,
                                                const long long* s_rowoff) {
  constexpr int kPer = EpiSlice<BN>::kPer;
  if (half * kPer >= BN) return;
  const uint32_t s = static_cast<uint32_t>(box_ctr) & 1;
  coop_res_issue(p, s_rowoff, n_t * BN + half * kPer, box_base + s * kRedBoxBytes, box_base + (2 + s) * kRedBoxBytes, r);
}

__device__ __forceinline__ void add_bf16x8_smem(uint32_t addr, float* v) {
  uint32_t w[4];
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(addr));
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[i]));
    v[2 * i] += f.x, v[2 * i + 1] += f.y;
  }
}

template <int BN>
__device__ __forceinline__ void epilogue_tile_coop(const GemmParams& p, int n_t, uint32_t t_acc, int r, int half,
                                                   const float* s_bias, const float* s_scale, uint32_t box_base, int& box_ctr,
                                                   const long long* s_rowoff) {
  constexpr int kPer = EpiSlice<BN>::kPer;
  const int cbeg = half * kPer;
  if (cbeg >= BN) return;
  uint32_t raw[32];
  __syncwarp();
  tmem_ld32(t_acc + cbeg, raw);
  const uint32_t sw = (static_cast<uint32_t>(r) >> 1) & 3;
#pragma unroll 1
  for (int c0 = cbeg; c0 < cbeg + kPer; c0 += 32) {
    const int col = n_t * BN + c0;
    const uint32_t s = static_cast<uint32_t>(box_ctr) & 1;
    const uint32_t box_a = box_base + s * kRedBoxBytes, box_b = box_base + (2 + s) * kRedBoxBytes;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    named_bar_sync(2 + half, 128);   // A: residual boxes of this box have landed; box k-1 has been copied out by everyone
    if (c0 + 32 < cbeg + kPer)
      coop_res_issue(p, s_rowoff, col + 32, box_base + (s ^ 1) * kRedBoxBytes, box_base + (2 + (s ^ 1)) * kRedBoxBytes, r);
    tmem_ld_wait();
    float v[32];
#pragma unroll
    for (int g = 0; g < 2; ++g) {
      float b[16];
      lds_f32x16(s_bias + c0 + 16 * g, b);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[16 * g + i] = __uint_as_float(raw[16 * g + i]) + b[i];
      if (p.act == ACT_GELU) {
#pragma unroll
        for (int i = 0; i < 16; i += 2) gelu_fast2(v[16 * g + i], v[16 * g + i + 1]);
      } else if (p.act == ACT_RELU) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[16 * g + i] = fmaxf(v[16 * g + i], 0.f);
      }
      if (p.scale) {
        float sc[16];
        lds_f32x16(s_scale + c0 + 16 * g, sc);
#pragma unroll
        for (int i = 0; i < 16; ++i) v[16 * g + i] *= sc[i];
      }
    }
    if (c0 + 32 < cbeg + kPer) {   // raw is dead: the next accumulator chunk travels behind the rest of this box
      __syncwarp();
      tmem_ld32(t_acc + c0 + 32, raw);
    }
    const uint32_t row_a = box_a + static_cast<uint32_t>(r) * 64, row_b = box_b + static_cast<uint32_t>(r) * 64;
    if (col < p.N) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const uint32_t o = (static_cast<uint32_t>(c) ^ sw) << 4;
        if (p.res1) add_bf16x8_smem(row_a + o, v + 8 * c);
        if (p.res2) add_bf16x8_smem(row_b + o, v + 8 * c);
      }
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const uint32_t o = (static_cast<uint32_t>(c) ^ sw) << 4;
      uint32_t w[4], wr[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float a0 = v[8 * c + 2 * i], a1 = v[8 * c + 2 * i + 1];
        __nv_bfloat162 h = __floats2bfloat162_rn(a0, a1);
        w[i] = *reinterpret_cast<uint32_t*>(&h);
        h = __floats2bfloat162_rn(fmaxf(a0, 0.f), fmaxf(a1, 0.f));
        wr[i] = *reinterpret_cast<uint32_t*>(&h);
      }
      asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_a + o), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
      if (p.out_relu)
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_b + o), "r"(wr[0]), "r"(wr[1]), "r"(wr[2]), "r"(wr[3]) : "memory");
    }
    named_bar_sync(2 + half, 128);   // B: the result boxes are complete
    if (col < p.N) {
      const uint32_t chunk = static_cast<uint32_t>(r) & 3;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t row = (static_cast<uint32_t>(r) >> 2) + 32 * i;
        const long long ro = s_rowoff[row];
        const uint32_t src = row * 64 + ((chunk ^ ((row >> 1) & 3)) << 4);
        if (ro >= 0) {
          uint4 t;
          if (p.out) {
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(t.x), "=r"(t.y), "=r"(t.z), "=r"(t.w) : "r"(box_a + src));
            *reinterpret_cast<uint4*>(p.out + ro + col + chunk * 8) = t;
          }
          if (p.out_relu) {
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(t.x), "=r"(t.y), "=r"(t.z), "=r"(t.w) : "r"(box_b + src));
            *reinterpret_cast<uint4*>(p.out_relu + ro + col + chunk * 8) = t;
          }
        }
      }
    }
    ++box_ctr;
  }
}

// t_acc: TMEM address (lane quadrant of this warp, accumulator stage, column 0 of the tile).
// s_bias / s_scale: staged bias / scale of the tile's BN columns (see stage_bias_scale).
template <int BN>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, int m_tile, int n_t, uint32_t t_acc, int r, int half,
                                              const float* s_bias, const float* s_scale, const TileSched& ts, ResRing& rr,
                                              int tile_iter) {
  const RowMap rm = map_row(p, m_tile, r);
  constexpr int kPer = EpiSlice<BN>::kPer;
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
  const bool staged = rr.nbuf > 0;
  const bool pre = staged && rm.valid;
  constexpr int kStep = 16;
  constexpr int kGroups = kPer / 16;
  const int step = (p.act == ACT_GEGLU) ? 32 : kStep;
#pragma unroll 1
  for (int c0 = cbeg; c0 < cbeg + kPer && c0 < BN; c0 += step) {
    const int col = n_t * BN + c0;
    uint32_t raw[16], gate[16];
    __syncwarp();
    tmem_ld16(t_acc + c0, raw);
    if (p.act == ACT_GEGLU) tmem_ld16(t_acc + c0 + 16, gate);
    float4 rpre[4];
    if (staged) {
      // request group k + nbuf - 1 (possibly of the next tile), then wait until group k has landed
      res_ring_issue<BN>(p, ts, rr, r, half);
      if (rr.nbuf == 4)
        asm volatile("cp.async.wait_group 3;" ::: "memory");
      else if (rr.nbuf == 3)
        asm volatile("cp.async.wait_group 2;" ::: "memory");
      else
        asm volatile("cp.async.wait_group 1;" ::: "memory");
      const int k = tile_iter * kGroups + ((c0 - cbeg) >> 4);
      const uint32_t slot = rr.base + static_cast<uint32_t>(k % rr.nbuf) * kResSlotPitch;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(rpre[i].x), "=f"(rpre[i].y), "=f"(rpre[i].z), "=f"(rpre[i].w)
                     : "r"(slot + 16 * i));
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
      for (int i = 0; i < 16; i += 2) {
        float g0 = __uint_as_float(gate[i]) + bg[i], g1 = __uint_as_float(gate[i + 1]) + bg[i + 1];
        gelu_fast2(g0, g1);
        v[i] = (__uint_as_float(raw[i]) + b[i]) * g0;
        v[i + 1] = (__uint_as_float(raw[i + 1]) + b[i + 1]) * g1;
      }
      if (p.split)
        store_split16(p.out + rm.row * p.ldo + (col >> 1), v, p.split);
      else
        store_bf16x16(p.out + rm.row * p.ldo + (col >> 1), v);
      continue;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(raw[i]) + b[i];
    finish_group16(p, rm, col, v, s_scale + c0, pre ? rpre : nullptr);  // unstaged residuals are loaded inside
  }
}

}  // namespace vda

// Fused depth tail on tcgen05 (dpt_temporal.py:116-123, dpt.py:117-124):
//
//   depth = relu( w1 . relu( conv3x3_{C->32}( bilinear_{align_corners}(x, (H, W)) ) + b3 ) + b1 )
//
// x: NHWC bf16 [n, h, w, C] (output_conv1's map, 296 x 296 x 128 for vitl), depth: f32 [n, H, W] (518 x 518).
// The two-kernel form writes the upsampled map (2.2 GB per vitl window) and reads it back: 5.1 GB of HBM traffic for
// 0.75 GB of algorithmic bytes.  Here the upsampled map never exists: per 8 x 16-pixel output tile
//   * the TMA producer fetches the source pixels under the tile's 10 x 18 halo patch (at most 8 x 12 of them),
//   * twenty interpolation warps build the UPSAMPLED halo patch straight in shared memory, one thread per (patch
//     column, 8 channels, quarter of the rows): the source row pair is reduced to its horizontal lerp in fp32 registers
//     (packed fp32x2 arithmetic) and every patch row is one vertical lerp, rounded to bf16 and stored in the
//     SWIZZLE_128B pattern that a TMA box would have produced (same arithmetic, to the bit, as bilinear_nhwc_kernel +
//     the conv's zero padding),
//   * the MMA warp reads the nine taps as tap-shifted UMMA descriptors into that patch (gemm.cu, halo mode) against
//     the 9 x C/64 weight tiles, which stay resident in shared memory for the whole kernel,
//   * four epilogue warps do bias + ReLU + the 32-wide dot + ReLU on the TMEM accumulator and store one float per pixel.
// Patches, source tiles and accumulators are double-buffered, so interpolation, MMAs and epilogue of consecutive
// tiles overlap.  HBM traffic: the source map once (plus halo overlap from L2) and the depth map.
#include <string.h>

#include "common.cuh"
#include "api_internal.h"
#include "../../include/vda_b200.h"

namespace vda {

constexpr int kTW = 8, kTH = 16;                 // output tile
constexpr int kPW = kTW + 2, kPH = kTH + 2;      // halo patch
constexpr int kSW = 8, kSH = 12;                 // source pixels under a patch (checked on the host)
constexpr int kPatchBytes = kPW * kPH * 128;     // per 64-channel block
constexpr int kPatchSlot = (kPatchBytes + 1023) / 1024 * 1024;
constexpr int kSrcBytes = kSW * kSH * 128;       // per 64-channel block
constexpr int kWTile = 32 * 128;                 // 32 output channels x 64 input channels
constexpr int kTailCtrlWarps = 4, kTailEpiWarps = 4;

struct alignas(64) TailParams {
  CUtensorMap tmX;   // 4D (C, w, h, n) bf16, box (64, kSW, kSH, 1), no swizzle
  CUtensorMap tmW;   // 2D (9*C, 32) bf16, box (64, 32), SWIZZLE_128B
  const float* b3;   // [32]
  const float* w1;   // [32]
  float b1;
  float* out;        // [n, H, W]
  int n, h, w, H, W;
  int tiles_w, tiles_h, num_tiles;
  float sy, sx;
};

template <int CB>
struct TailCfg {
  static constexpr int kRowParts = 4;           // the 18 patch rows are split 5 + 5 + 4 + 4 over four threads
  static constexpr int kInterpWarps = 10 * CB;  // 10 patch columns x 8*CB channel chunks x 4 row parts = 320*CB threads
  static constexpr int kThreads = 32 * (kTailCtrlWarps + kTailEpiWarps + kInterpWarps);
  static constexpr int kWBytes = 9 * CB * kWTile;
  static constexpr int kSrcStage = CB * kSrcBytes;
  static constexpr int kPatchStage = CB * kPatchSlot;
  static constexpr int kSmem = 1024 + kWBytes + 2 * kPatchStage + 2 * kSrcStage + 512;
};

__device__ __forceinline__ void tail_tile_coord(const TailParams& p, int tile, int& img, int& Y0, int& X0) {
  const int tpi = p.tiles_w * p.tiles_h;
  img = tile / tpi;
  const int t = tile - img * tpi;
  const int ty = t / p.tiles_w;
  Y0 = ty * kTH;
  X0 = (t - ty * p.tiles_w) * kTW;
}
// first source row / column needed by a patch that starts at output coordinate c0 - 1
__device__ __forceinline__ int tail_src_base(float scale, int c0, int limit) {
  const int c = c0 > 0 ? c0 - 1 : 0;
  return min(static_cast<int>(scale * c), limit - 1);
}

template <int CB>
__global__ void __launch_bounds__(TailCfg<CB>::kThreads, 1) depth_tail_tc_kernel(const __grid_constant__ TailParams p) {
  using Cfg = TailCfg<CB>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* sW = smem;                                   // 9*CB weight tiles, resident
  uint8_t* sPatch = sW + Cfg::kWBytes;                  // [2][CB][kPatchSlot]
  uint8_t* sSrc = sPatch + 2 * Cfg::kPatchStage;        // [2][CB][kSrcBytes]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sSrc + 2 * Cfg::kSrcStage);
  uint64_t* w_full = bars;
  uint64_t* src_full = bars + 1;     // [2]
  uint64_t* src_empty = bars + 3;    // [2]
  uint64_t* patch_full = bars + 5;   // [2]
  uint64_t* patch_empty = bars + 7;  // [2]
  uint64_t* tfull = bars + 9;        // [2]
  uint64_t* tempty = bars + 11;      // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);
  float* s_b3 = reinterpret_cast<float*>(bars + 16);    // 32 + 32 floats
  float* s_w1 = s_b3 + 32;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmX);
    tma_prefetch_desc(&p.tmW);
    mbar_init(w_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&src_full[s], 1);
      mbar_init(&src_empty[s], Cfg::kInterpWarps);
      mbar_init(&patch_full[s], Cfg::kInterpWarps);
      mbar_init(&patch_empty[s], 1);
      mbar_init(&tfull[s], 1);
      mbar_init(&tempty[s], kTailEpiWarps);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 64);
    tmem_relinquish();
  }
  if (warp == 3) s_b3[lane] = p.b3[lane], s_w1[lane] = p.w1[lane];   // weights: never written by a preceding kernel
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_sync();  // activations from here on

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      mbar_expect_tx(w_full, Cfg::kWBytes);
      for (int i = 0; i < 9 * CB; ++i) tma_load_2d(sW + i * kWTile, &p.tmW, w_full, i * 64, 0);
    }
    __syncwarp();
    int it = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
      const int s = it & 1;
      int img, Y0, X0;
      tail_tile_coord(p, tile, img, Y0, X0);
      const int sy0 = tail_src_base(p.sy, Y0, p.h), sx0 = tail_src_base(p.sx, X0, p.w);
      mbar_wait_parked(&src_empty[s], ((it >> 1) & 1) ^ 1);
      if (elect_one()) {
        mbar_expect_tx(&src_full[s], Cfg::kSrcStage);
        for (int cb = 0; cb < CB; ++cb)
          tma_load_4d(sSrc + s * Cfg::kSrcStage + cb * kSrcBytes, &p.tmX, &src_full[s], cb * 64, sx0, sy0, img);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = umma_idesc_f16(128, 32, true);
    mbar_wait(w_full, 0);
    int it = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
      const int s = it & 1;
      const uint32_t ph = (it >> 1) & 1;
      mbar_wait_parked(&tempty[s], ph ^ 1);
      mbar_wait_parked(&patch_full[s], ph);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + s * 32;
      const uint32_t patch = smem_u32(sPatch + s * Cfg::kPatchStage);
      const uint32_t wbase = smem_u32(sW);
      if (elect_one()) {
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          const int dy = tap / 3, dx = tap - dy * 3;
#pragma unroll
          for (int cb = 0; cb < CB; ++cb) {
            const uint64_t adesc = umma_desc_k_sw128_sbo(patch + cb * kPatchSlot + (dy * kPW + dx) * 128, kPW * 128);
            const uint64_t bdesc = umma_desc_k_sw128(wbase + (tap * CB + cb) * kWTile);
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (tap | cb | k) != 0);
          }
        }
        umma_commit(&patch_empty[s]);
        umma_commit(&tfull[s]);
      }
      __syncwarp();
    }
  } else if (warp >= kTailCtrlWarps && warp < kTailCtrlWarps + kTailEpiWarps) {
    // ------------------------------------------------------------------ epilogue: row r = pixel (r >> 3, r & 7) of the tile
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
      const int s = it & 1;
      int img, Y0, X0;
      tail_tile_coord(p, tile, img, Y0, X0);
      mbar_wait_parked(&tfull[s], (it >> 1) & 1);
      tc_fence_after();
      uint32_t raw[32];
      tmem_ld32(tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + s * 32, raw);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[s]);
      float dot = p.b1;
#pragma unroll
      for (int i = 0; i < 32; ++i) dot += fmaxf(__uint_as_float(raw[i]) + s_b3[i], 0.f) * s_w1[i];
      const int Y = Y0 + (r >> 3), X = X0 + (r & 7);
      if (Y < p.H && X < p.W) p.out[(static_cast<long long>(img) * p.H + Y) * p.W + X] = fmaxf(dot, 0.f);
    }
  } else if (warp >= kTailCtrlWarps + kTailEpiWarps) {
    // ------------------------------------------------------------------ interpolation: one (patch column, 8 channels,
    // row half) per thread; arithmetic identical to bilinear_nhwc_kernel (elementwise.cu)
    const int t = threadIdx.x - 32 * (kTailCtrlWarps + kTailEpiWarps);
    constexpr int kChunks = 8 * CB;
    const int chunk = t % kChunks;
    const int px = (t / kChunks) % kPW;
    const int part = t / (kChunks * kPW);
    const int cb = chunk >> 3, k = chunk & 7;
    const int py_begin = part < 2 ? 5 * part : 10 + 4 * (part - 2), py_end = py_begin + (part < 2 ? 5 : 4);
    int it = 0;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
      const int s = it & 1;
      const uint32_t ph = (it >> 1) & 1;
      int img, Y0, X0;
      tail_tile_coord(p, tile, img, Y0, X0);
      const int sy0 = tail_src_base(p.sy, Y0, p.h), sx0 = tail_src_base(p.sx, X0, p.w);
      const int X = X0 - 1 + px;
      const bool col_in = X >= 0 && X < p.W;
      const float fx = p.sx * X;
      const int x0 = min(static_cast<int>(fx), p.w - 1);
      const int x1 = min(x0 + 1, p.w - 1);
      const float lx = fx - x0;
      const uint64_t lx2 = f2pack(lx, lx);
      mbar_wait_parked(&src_full[s], ph, 200);
      mbar_wait_parked(&patch_empty[s], ph ^ 1, 200);
      const uint32_t src = smem_u32(sSrc + s * Cfg::kSrcStage + cb * kSrcBytes) + k * 16;
      const uint32_t o0 = static_cast<uint32_t>(x0 - sx0) * 128, o1 = static_cast<uint32_t>(x1 - sx0) * 128;
      const uint32_t dst = smem_u32(sPatch + s * Cfg::kPatchStage + cb * kPatchSlot);
      // hA / hB: horizontal lerps (8 channels as 4 fp32 pairs) of the two source rows under the current patch row.
      // Which of the two holds the upper row alternates (`flip`), so moving down by one source row reloads one
      // buffer and copies nothing.
      uint64_t hA[4], hB[4];
      auto load_row = [&](int y, uint64_t (&hh)[4]) {
        const uint32_t row = src + static_cast<uint32_t>(y - sy0) * (kSW * 128);
        uint32_t pp[4], qq[4];
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(pp[0]), "=r"(pp[1]), "=r"(pp[2]), "=r"(pp[3]) : "r"(row + o0));
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(qq[0]), "=r"(qq[1]), "=r"(qq[2]), "=r"(qq[3]) : "r"(row + o1));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          // bf16x2 word -> (fp32, fp32): low half shifted up, high half masked; fq - fp and the fma as in the
          // scalar kernel (fmaf(lx, fq - fp, fp)), two channels per instruction
          const uint64_t fp = f2pack(__uint_as_float(pp[i] << 16), __uint_as_float(pp[i] & 0xffff0000u));
          const uint64_t fq = f2pack(__uint_as_float(qq[i] << 16), __uint_as_float(qq[i] & 0xffff0000u));
          hh[i] = f2fma(lx2, f2sub(fq, fp), fp);
        }
      };
      auto emit = [&](uint32_t at, float ly, const uint64_t (&lo)[4], const uint64_t (&hi)[4]) {
        const uint64_t ly2 = f2pack(ly, ly);
        uint32_t o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float r0, r1;
          f2unpack(f2fma(ly2, f2sub(hi[i], lo[i]), lo[i]), r0, r1);
          const __nv_bfloat162 r2 = __floats2bfloat162_rn(r0, r1);
          o[i] = *reinterpret_cast<const uint32_t*>(&r2);
        }
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(at), "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
      };
      int ya = -1, yb = -1;   // source rows held by hA / hB
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        const int py = py_begin + i;
        if (py < py_end) {
          const int Y = Y0 - 1 + py;
          const int R = py * kPW + px;   // patch row
          const uint32_t at = dst + static_cast<uint32_t>(R) * 128 + (static_cast<uint32_t>(k ^ (R & 7)) << 4);
          if (col_in && Y >= 0 && Y < p.H) {
            const float fy = p.sy * Y;
            const int y0 = min(static_cast<int>(fy), p.h - 1);
            const int y1 = min(y0 + 1, p.h - 1);
            const float ly = fy - y0;
            if (y0 == ya) {            // hA is the upper row
              if (y1 != yb) load_row(y1, hB), yb = y1;
              emit(at, ly, hA, hB);
            } else if (y0 == yb) {     // moved down by one source row: hB becomes the upper row
              if (y1 != ya) load_row(y1, hA), ya = y1;
              emit(at, ly, hB, hA);
            } else {
              load_row(y0, hA), ya = y0;
              if (y1 != yb) load_row(y1, hB), yb = y1;
              emit(at, ly, hA, hB);
            }
          } else {                     // outside the image: the convolution's zero padding
            asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(at), "r"(0u) : "memory");
          }
        }
      }
      fence_proxy_async_smem();   // generic-proxy writes of the patch -> visible to the UMMA (async proxy) reads
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&patch_full[s]);
        mbar_arrive(&src_empty[s]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

template <int CB>
static int launch_tail(const TailParams& p, cudaStream_t st) {
  using Cfg = TailCfg<CB>;
  static DeviceOnce once;
  int& configured = once.slot();
  if (!configured) {
    int rc = check_cuda(cudaFuncSetAttribute(depth_tail_tc_kernel<CB>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmem),
                        "depth_tail smem attribute");
    if (rc) return rc;
    configured = 1;
  }
  const int grid = p.num_tiles < num_sms() ? p.num_tiles : num_sms();
  return check_cuda(launch_pdl(depth_tail_tc_kernel<CB>, dim3(grid), dim3(Cfg::kThreads), Cfg::kSmem, st, p), "depth_tail launch");
}

}  // namespace vda

using namespace vda;

extern "C" int vda_depth_tail(const void* x, int ldx, const void* w3, const float* b3, const float* w1, float b1, float* out,
                              int n, int h, int w, int H, int W, int C, void* stream) {
  if (!x || !w3 || !b3 || !w1 || !out || n <= 0 || h <= 1 || w <= 1 || H < h || W < w || ldx % 8 || ldx < C)
    return set_error(-1, "vda_depth_tail: bad arguments");
  if (C != 64 && C != 128) return set_error(-1, "vda_depth_tail: C=%d unsupported (64 or 128)", C);
  if (num_sms() <= 0) return set_error(-2, "vda_init(device) has not been called for the current device");
  TailParams p;
  memset(&p, 0, sizeof(p));
  p.sy = static_cast<float>(h - 1) / (H - 1);
  p.sx = static_cast<float>(w - 1) / (W - 1);
  // source pixels under one 10 x 18 halo patch: floor((rows - 1) * scale) + 1 between the first taps of the first
  // and the last row, + 1 for the second tap, + 1 to turn a difference into a count
  // (floor(a + d) - floor(a) <= floor(d) + 1)
  if (static_cast<int>((kPH - 1) * static_cast<double>(p.sy) + 1e-3) + 3 > kSH ||
      static_cast<int>((kPW - 1) * static_cast<double>(p.sx) + 1e-3) + 3 > kSW)
    return set_error(-7, "vda_depth_tail: upsampling ratio %dx%d -> %dx%d needs more than %d x %d source pixels per patch", h, w,
                     H, W, kSH, kSW);
  {
    const uint64_t dims[4] = {(uint64_t)C, (uint64_t)w, (uint64_t)h, (uint64_t)n};
    const uint64_t str[3] = {(uint64_t)ldx * 2, (uint64_t)w * ldx * 2, (uint64_t)h * w * ldx * 2};
    const uint32_t box[4] = {64, kSW, kSH, 1};
    int rc = encode_bf16_map_plain(&p.tmX, x, 4, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)(9 * C), 32};
    const uint64_t str[1] = {(uint64_t)(9 * C) * 2};
    const uint32_t box[2] = {64, 32};
    int rc = encode_bf16_map(&p.tmW, w3, 2, dims, str, box);
    if (rc) return rc;
  }
  p.b3 = b3, p.w1 = w1, p.b1 = b1, p.out = out;
  p.n = n, p.h = h, p.w = w, p.H = H, p.W = W;
  p.tiles_w = (W + kTW - 1) / kTW, p.tiles_h = (H + kTH - 1) / kTH;
  p.num_tiles = n * p.tiles_w * p.tiles_h;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return C == 128 ? launch_tail<2>(p, st) : launch_tail<1>(p, st);
}

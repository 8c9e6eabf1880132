// Fused depth tail (dpt_temporal.py:116-123, dpt.py:117-124):
//   bilinear(align_corners=True) [n,h,w,C] -> [n,H,W,C]  ->  3x3 conv C->32 + bias + ReLU  ->  1x1 conv 32->1 + bias + ReLU
// in one kernel, so that the upsampled [n,H,W,C] tensor (2.2 GB per vitl window) is never written and
// the 3x3 taps re-use an upsampled halo patch held in shared memory.  Unfused, the pair costs
// 0.75 ms (upsample, HBM-bound) + 2.5 ms (Cout = 32 implicit GEMM: every A tile is fetched 9 times
// through TMA and the kernel runs at the L2->SM bandwidth, ~8 TB/s).
//
// One CTA per SM walks 16x16-pixel output tiles:
//   phase 1  all 256 threads interpolate the 18x18 halo patch (zero outside the image = conv padding),
//            same arithmetic and bf16 rounding as bilinear_nhwc_kernel, into padded smem rows;
//   phase 2  implicit GEMM M = 256 pixels, N = 32, K = 9*C on mma.sync.m16n8k16: A fragments are
//            ldmatrix reads of the patch at tap-shifted pixel rows (per-lane row addresses make the
//            shift free), B fragments come from the conv weights staged once per CTA.  Warp (wm, kh)
//            owns tile rows 4wm..4wm+3 and the channel half kh of every tap (split-K), so each
//            (pixel, k) element of the patch is read from smem exactly once per tap;
//   phase 3  the two channel halves are summed through smem; bias + ReLU, the 32-wide dot with the
//            1x1 weights, bias + ReLU, fp32 store.
// Bound: shared-memory bandwidth (A traffic = 9 * 256 * C * 2 B per tile); the tensor work is small.
//
// Status (measured on B200, vitl tail 32 x 296^2 x 128 -> 518^2): 3.30 ms against 3.25 ms for the
// two-kernel path, so the engine keeps the two kernels by default (VDA_TAIL_FUSED=1 selects this one).
// Phase 1 alone takes 1.9 ms: with 194 KB of shared memory carved out the L1 is too small to hold the
// source footprint, the 4 taps x 3 overlapping outputs re-read it from L2 (11.6 GB per window, i.e.
// the same L2 bound as the unfused conv), and more warps do not help (16 warps: 2.25 ms).  Phase 2
// alone takes 1.5 ms (mma.sync + ldmatrix at 73 B/clk of shared-memory traffic).  The next step is a
// TMA-staged source patch (each source byte fetched once) and a tcgen05 MMA on tap-shifted descriptors.
#include <stdlib.h>

#include "common.cuh"
#include "api_internal.h"
#include "../../include/vda_b200.h"

namespace vda {

struct TailParams {
  const bf16* x;
  int ldx;
  const bf16* w3;   // [32][9*C], k = (kh, kw, c)
  const float* b3;  // [32]
  const float* w1;  // [32]
  float b1;
  float* out;       // [n, H, W]
  int n, h, w, H, W;
  float sy, sx;
  int tiles_x, tiles_y, num_tiles;
};

constexpr int kTT = 16;         // tile edge (output pixels)
constexpr int kTH = kTT + 2;    // halo edge

__device__ __forceinline__ void tl_ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void tl_mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t tl_mix2(uint32_t a, uint32_t b, uint32_t c, uint32_t d, float w00, float w01, float w10,
                                            float w11) {
  const float2 fa = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&a));
  const float2 fb = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&b));
  const float2 fc = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&c));
  const float2 fd = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&d));
  const __nv_bfloat162 r = __floats2bfloat162_rn(w00 * fa.x + w01 * fb.x + w10 * fc.x + w11 * fd.x,
                                                 w00 * fa.y + w01 * fb.y + w10 * fc.y + w11 * fd.y);
  return *reinterpret_cast<const uint32_t*>(&r);
}

template <int C, int DBG = 0>  // DBG (timing experiments): 1 = skip the MMA phase, 2 = skip the interpolation phase
__global__ void __launch_bounds__(256, 1) depth_tail_kernel(const TailParams p) {
  constexpr int kPS = C * 2 + 16;       // patch row pitch (bytes): +16 keeps ldmatrix conflict-free
  constexpr int kWS = 9 * C * 2 + 16;   // weight row pitch
  constexpr int kC8 = C / 8;
  extern __shared__ __align__(16) uint8_t tail_sm[];
  uint8_t* sW = tail_sm;
  uint8_t* sP = tail_sm + 32 * kWS;
  const uint32_t sW_u = smem_u32(sW), sP_u = smem_u32(sP);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  for (int i = tid; i < 32 * (9 * kC8); i += 256) {
    const int n = i / (9 * kC8), c = i - n * (9 * kC8);
    *reinterpret_cast<uint4*>(sW + n * kWS + c * 16) = __ldg(reinterpret_cast<const uint4*>(p.w3 + static_cast<long long>(n) * 9 * C) + c);
  }
  float bias[8], wdot[8];  // this lane's 8 output channels: nt*8 + (lane%4)*2 + j
#pragma unroll
  for (int nt = 0; nt < 4; ++nt)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int col = nt * 8 + (lane & 3) * 2 + j;
      bias[nt * 2 + j] = __ldg(p.b3 + col);
      wdot[nt * 2 + j] = __ldg(p.w1 + col);
    }
  __syncthreads();

  const int wm = warp & 3, kh = warp >> 2;
  for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
    const int per_img = p.tiles_x * p.tiles_y;
    const int img = tile / per_img;
    const int rem = tile - img * per_img;
    const int ty = rem / p.tiles_x, tx = rem - ty * p.tiles_x;
    const int Y0 = ty * kTT, X0 = tx * kTT;
    const bf16* base = p.x + static_cast<long long>(img) * p.h * p.w * p.ldx;

    // ---- phase 1: upsampled halo patch.  kU units per thread are in flight together (4*kU 16-byte
    // loads) because with one CTA of 8 warps per SM this phase is bound by the L2 load latency.
    constexpr int kUnits = kTH * kTH * kC8;
    constexpr int kU = 5;
#pragma unroll 1
    for (int i0 = 0; i0 < kUnits; i0 += 256 * kU) {
      uint4 ta[kU], tb[kU], tc[kU], td[kU];
      float w00[kU], w01[kU], w10[kU], w11[kU];
      bool in[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int idx = i0 + u * 256 + tid;
        const int px = idx / kC8, ch = idx - px * kC8;
        const int py = px / kTH, pxx = px - py * kTH;
        const int Y = Y0 - 1 + py, X = X0 - 1 + pxx;
        in[u] = DBG != 2 && idx < kUnits && Y >= 0 && Y < p.H && X >= 0 && X < p.W;
        ta[u] = tb[u] = tc[u] = td[u] = make_uint4(0, 0, 0, 0);
        if (in[u]) {
          const float fy = p.sy * Y, fx = p.sx * X;
          int y0 = static_cast<int>(fy), x0 = static_cast<int>(fx);
          y0 = min(y0, p.h - 1), x0 = min(x0, p.w - 1);
          const int y1 = min(y0 + 1, p.h - 1), x1 = min(x0 + 1, p.w - 1);
          const float ly = fy - y0, lx = fx - x0;
          w00[u] = (1.f - ly) * (1.f - lx), w01[u] = (1.f - ly) * lx, w10[u] = ly * (1.f - lx), w11[u] = ly * lx;
          const bf16* c0 = base + ch * 8;
          ta[u] = __ldg(reinterpret_cast<const uint4*>(c0 + (static_cast<long long>(y0) * p.w + x0) * p.ldx));
          tb[u] = __ldg(reinterpret_cast<const uint4*>(c0 + (static_cast<long long>(y0) * p.w + x1) * p.ldx));
          tc[u] = __ldg(reinterpret_cast<const uint4*>(c0 + (static_cast<long long>(y1) * p.w + x0) * p.ldx));
          td[u] = __ldg(reinterpret_cast<const uint4*>(c0 + (static_cast<long long>(y1) * p.w + x1) * p.ldx));
        }
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int idx = i0 + u * 256 + tid;
        if (idx < kUnits) {
          const int px = idx / kC8, ch = idx - px * kC8;
          uint4 v = make_uint4(0, 0, 0, 0);
          if (in[u]) {
            v.x = tl_mix2(ta[u].x, tb[u].x, tc[u].x, td[u].x, w00[u], w01[u], w10[u], w11[u]);
            v.y = tl_mix2(ta[u].y, tb[u].y, tc[u].y, td[u].y, w00[u], w01[u], w10[u], w11[u]);
            v.z = tl_mix2(ta[u].z, tb[u].z, tc[u].z, td[u].z, w00[u], w01[u], w10[u], w11[u]);
            v.w = tl_mix2(ta[u].w, tb[u].w, tc[u].w, td[u].w, w00[u], w01[u], w10[u], w11[u]);
          }
          *reinterpret_cast<uint4*>(sP + px * kPS + ch * 16) = v;
        }
      }
    }
    __syncthreads();

    // ---- phase 2: implicit GEMM on the patch
    float acc[4][4][4];
#pragma unroll
    for (int mt = 0; mt < 4; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[mt][nt][i] = 0.f;
#pragma unroll 1
    for (int tap = 0; tap < (DBG == 1 ? 0 : 9); ++tap) {
      const int dy = tap / 3, dx = tap - dy * 3;
      // lane's A row: tile column (lane & 15) shifted by the tap, 16-byte half (lane >> 4) of the k16 step
      const uint32_t a_lane = sP_u + ((wm * 4 + dy) * kTH + (lane & 15) + dx) * kPS + (lane >> 4) * 16;
      const uint32_t b_lane = sW_u + ((lane & 7) + ((lane >> 4) << 3)) * kWS + tap * C * 2 + ((lane >> 3) & 1) * 16;
#pragma unroll
      for (int kc = 0; kc < C / 32; ++kc) {
        const int ch0 = kh * (C / 2) + kc * 16;
        uint32_t b[2][4];
        tl_ldsm_x4(b_lane + ch0 * 2, b[0]);
        tl_ldsm_x4(b_lane + 16 * kWS + ch0 * 2, b[1]);
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) {
          uint32_t a[4];
          tl_ldsm_x4(a_lane + mt * kTH * kPS + ch0 * 2, a);
#pragma unroll
          for (int nt = 0; nt < 4; ++nt) tl_mma(acc[mt][nt], a, b[nt >> 1][(nt & 1) * 2], b[nt >> 1][(nt & 1) * 2 + 1]);
        }
      }
    }
    __syncthreads();  // every warp is done reading the patch: it becomes the exchange buffer

    // ---- phase 3: add the two channel halves, epilogue
    float* ex = reinterpret_cast<float*>(sP) + wm * (64 * 32) + lane;
    if (kh == 1) {
#pragma unroll
      for (int mt = 0; mt < 4; ++mt)
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
          for (int i = 0; i < 4; ++i) ex[((mt * 4 + nt) * 4 + i) * 32] = acc[mt][nt][i];
    }
    __syncthreads();
    if (kh == 0) {
#pragma unroll
      for (int mt = 0; mt < 4; ++mt) {
        const int Y = Y0 + wm * 4 + mt;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          float part = 0.f;
#pragma unroll
          for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const float v = acc[mt][nt][2 * hh + j] + ex[((mt * 4 + nt) * 4 + 2 * hh + j) * 32] + bias[nt * 2 + j];
              part = fmaf(fmaxf(v, 0.f), wdot[nt * 2 + j], part);
            }
          part += __shfl_xor_sync(0xffffffffu, part, 1);
          part += __shfl_xor_sync(0xffffffffu, part, 2);
          const int X = X0 + (lane >> 2) + 8 * hh;
          if ((lane & 3) == 0 && Y < p.H && X < p.W)
            p.out[(static_cast<long long>(img) * p.H + Y) * p.W + X] = fmaxf(part + p.b1, 0.f);
        }
      }
    }
    __syncthreads();  // exchange buffer is free again before the next patch is written
  }
}

template <int C>
static int launch_tail(const TailParams& p, int num_sms, cudaStream_t st) {
  constexpr int smem = 32 * (9 * C * 2 + 16) + kTH * kTH * (C * 2 + 16);
  static_assert(kTH * kTH * (C * 2 + 16) >= 4 * 64 * 32 * 4, "exchange buffer must fit in the patch");
  static DeviceOnce once;
  int& configured = once.slot();
  if (!configured) {
    int rc = check_cuda(cudaFuncSetAttribute(depth_tail_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                        "depth_tail smem attribute");
    if (rc) return rc;
    configured = 1;
  }
  const int grid = p.num_tiles < num_sms ? p.num_tiles : num_sms;
  const char* dbg = getenv("VDA_TAIL_DEBUG");
  if (dbg && atoi(dbg) == 1) {
    cudaFuncSetAttribute(depth_tail_kernel<C, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    depth_tail_kernel<C, 1><<<grid, 256, smem, st>>>(p);
  } else if (dbg && atoi(dbg) == 2) {
    cudaFuncSetAttribute(depth_tail_kernel<C, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    depth_tail_kernel<C, 2><<<grid, 256, smem, st>>>(p);
  } else
    depth_tail_kernel<C><<<grid, 256, smem, st>>>(p);
  return check_cuda(cudaGetLastError(), "depth_tail launch");
}

}  // namespace vda

using namespace vda;

extern "C" int vda_depth_tail(const void* x, int ldx, const void* w3, const float* b3, const float* w1, float b1, float* out,
                              int n, int h, int w, int H, int W, int C, void* stream) {
  if (!x || !w3 || !b3 || !w1 || !out || n <= 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0 || ldx % 8 || ldx < C)
    return set_error(-1, "vda_depth_tail: bad arguments");
  TailParams p;
  p.x = static_cast<const bf16*>(x), p.ldx = ldx, p.w3 = static_cast<const bf16*>(w3), p.b3 = b3, p.w1 = w1, p.b1 = b1;
  p.out = out, p.n = n, p.h = h, p.w = w, p.H = H, p.W = W;
  p.sy = H > 1 ? static_cast<float>(h - 1) / (H - 1) : 0.f;
  p.sx = W > 1 ? static_cast<float>(w - 1) / (W - 1) : 0.f;
  p.tiles_x = (W + kTT - 1) / kTT, p.tiles_y = (H + kTT - 1) / kTT;
  p.num_tiles = n * p.tiles_x * p.tiles_y;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (C == 128) return launch_tail<128>(p, num_sms(), st);
  if (C == 64) return launch_tail<64>(p, num_sms(), st);
  return set_error(-1, "vda_depth_tail: C=%d unsupported (64 or 128)", C);
}

// tcgen05 / TMEM / TMA GEMM and 3x3 implicit-GEMM kernel for sm_100a.
//
//   C[M,N] = epilogue( A[M,K] * W[N,K]^T )      A, W bf16 (K-major), fp32 accumulate in TMEM
//
// Replaces every nn.Linear / 1x1 conv / 3x3 conv / ConvTranspose(k=s) call site of the reference
// hot path: dinov2_layers/attention.py:51,60, mlp.py:36-39, dpt.py:60-124,
// util/blocks.py:78-91,160, motion_module/motion_module.py:115,128, motion_module/attention.py:81-83.
//
// One persistent CTA per SM, 12 warps:
//   warp 0   TMA producer   (A tile 128x64, W tile BLOCK_Nx64 per stage, 128B swizzle)
//   warp 1   MMA issuer     (one thread, tcgen05.mma cta_group::1 M=128 N=BLOCK_N K=16)
//   warp 2   TMEM allocator (2 accumulator stages of BLOCK_N fp32 columns)
//   warp 4-11 epilogue      (tcgen05.ld 32x32b -> registers -> fused epilogue -> global)
#include "common.cuh"
#include "gemm_params.h"

namespace vda {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;
constexpr int kGemmThreads = 384;
constexpr int kEpiWarps = 8;

template <int BN>
struct GemmCfg {
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = BN * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN >= 256) ? 4 : (BN >= 128 ? 6 : 8);
  static constexpr int kTmemCols = (2 * BN <= 32) ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

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
    rm.valid = (h < p.H) && (w < p.W);
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

template <int BN>
__global__ void __launch_bounds__(kGemmThreads, 1) gemm_tc_kernel(const __grid_constant__ GemmParams p) {
  using Cfg = GemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.m_tiles * p.n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], kEpiWarps);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const int cblocks = (p.a_mode == A_CONV3X3) ? (p.C / kBlockK) : 1;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_t = tile / p.n_tiles, n_t = tile - m_t * p.n_tiles;
        int img = 0, h0 = 0, w0 = 0;
        if (p.a_mode == A_CONV3X3) {
          const int tpi = p.tiles_w * p.tiles_h;
          img = m_t / tpi;
          const int t = m_t - img * tpi;
          const int ty = t / p.tiles_w, tx = t - ty * p.tiles_w;
          h0 = ty * (kBlockM >> p.tw_log2);
          w0 = tx << p.tw_log2;
        }
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kABytes;
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          if (p.a_mode == A_CONV3X3) {
            const int tap = kb / cblocks, cb = kb - tap * cblocks;
            const int kh = tap / 3, kw = tap - kh * 3;
            tma_load_4d(sa, &p.tmA, &full_bar[stage], cb * kBlockK, w0 + kw - 1, h0 + kh - 1, img);
          } else {
            tma_load_2d(sa, &p.tmA, &full_bar[stage], kb * kBlockK, m_t * kBlockM);
          }
          tma_load_2d(sb, &p.tmB, &full_bar[stage], kb * kBlockK, n_t * BN);
          if (++stage == Cfg::kStages) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_f16(kBlockM, BN, true);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tempty_bar[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint64_t adesc = umma_desc_k_sw128(sa);
          const uint64_t bdesc = umma_desc_k_sw128(sa + Cfg::kABytes);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            // +32 B per K=16 step inside the 128 B swizzle row (start-address field is >>4)
            umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          }
          umma_commit(&empty_bar[stage]);
          if (++stage == Cfg::kStages) stage = 0, phase ^= 1;
        }
        umma_commit(&tfull_bar[as]);
      }
    }
  } else if (warp >= 4) {
    // ---------------------------------------------------------------- epilogue
    const int ew = warp - 4;
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    const int half = ew >> 2;
    const int r = quad * 32 + lane;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int m_t = tile / p.n_tiles, n_t = tile - m_t * p.n_tiles;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      const RowMap rm = map_row(p, m_t, r);
      mbar_wait(&tfull_bar[as], aphase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + as * BN;

      if (p.store_mode == ST_HEADDOT) {
        if (half == 0) {
          float dot = p.b2;
#pragma unroll
          for (int c0 = 0; c0 < BN && c0 < 32; c0 += 16) {
            uint32_t raw[16];
            tmem_ld16(t_row + c0, raw);
            tmem_ld_wait();
            if (c0 < p.N) {
              float b[16], w[16];
              load_f32x16(p.bias + c0, b);
              load_f32x16(p.w2 + c0, w);
#pragma unroll
              for (int i = 0; i < 16; ++i) dot += fmaxf(__uint_as_float(raw[i]) + b[i], 0.f) * w[i];
            }
          }
          if (rm.valid) p.out_f32[rm.row] = fmaxf(dot, 0.f);
        }
      } else if (p.act == ACT_GEGLU) {
        // physical columns: [32g,32g+16) value, [32g+16,32g+32) gate -> output cols [16g,16g+16)
        constexpr int kPer = (BN / 2 >= 32) ? BN / 2 : 32;
        const int cbeg = half * kPer;
        for (int c0 = cbeg; c0 < cbeg + kPer && c0 < BN; c0 += 32) {
          uint32_t rv[16], rg[16];
          __syncwarp();
          tmem_ld16(t_row + c0, rv);
          tmem_ld16(t_row + c0 + 16, rg);
          tmem_ld_wait();
          const int col = n_t * BN + c0;
          if (col < p.N && rm.valid) {
            float bv[16], bg[16], v[16];
            load_f32x16(p.bias + col, bv);
            load_f32x16(p.bias + col + 16, bg);
#pragma unroll
            for (int i = 0; i < 16; ++i)
              v[i] = (__uint_as_float(rv[i]) + bv[i]) * gelu_erf(__uint_as_float(rg[i]) + bg[i]);
            store_bf16x16(p.out + rm.row * p.ldo + (col >> 1), v);
          }
        }
      } else {
        constexpr int kPer = (BN / 2 >= 16) ? BN / 2 : 16;
        const int cbeg = half * kPer;
        for (int c0 = cbeg; c0 < cbeg + kPer && c0 < BN; c0 += 16) {
          uint32_t raw[16];
          __syncwarp();
          tmem_ld16(t_row + c0, raw);
          tmem_ld_wait();
          const int col = n_t * BN + c0;
          if (col >= p.N || !rm.valid) continue;
          float v[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(raw[i]);
          if (p.bias) {
            float b[16];
            load_f32x16(p.bias + col, b);
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] += b[i];
          }
          if (p.act == ACT_GELU) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = gelu_erf(v[i]);
          } else if (p.act == ACT_RELU) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
          }
          if (p.scale) {
            float s[16];
            load_f32x16(p.scale + col, s);
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] *= s[i];
          }
          long long off;
          if (p.store_mode == ST_CONVT) {
            const int g = col / p.ct_cout, co = col - g * p.ct_cout;
            const int kh = g / p.ct_s, kw = g - kh * p.ct_s;
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
            const int b = m / p.qkv_s, s = m - b * p.qkv_s;
            const int cc = col - 2 * p.qkv_d;
            const int head = cc >> 6, d0 = cc & 63;
            bf16* dst = p.vt + (static_cast<long long>(b * p.qkv_heads + head) * 64 + d0) * p.qkv_spad + s;
#pragma unroll
            for (int i = 0; i < 16; ++i) dst[static_cast<long long>(i) * p.qkv_spad] = __float2bfloat16(v[i]);
            continue;
          }
          if (p.res_f32) {
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
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

template <int BN>
static cudaError_t launch_bn(const GemmParams& p, int num_sms, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  const int tiles = p.m_tiles * p.n_tiles;
  const int grid = tiles < num_sms ? tiles : num_sms;
  gemm_tc_kernel<BN><<<grid, kGemmThreads, Cfg::kSmemBytes, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_gemm(const GemmParams& p, int block_n, int num_sms, cudaStream_t stream) {
  switch (block_n) {
    case 32: return launch_bn<32>(p, num_sms, stream);
    case 64: return launch_bn<64>(p, num_sms, stream);
    case 128: return launch_bn<128>(p, num_sms, stream);
    case 256: return launch_bn<256>(p, num_sms, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace vda

// Probe of the tcgen05 shared-memory descriptor semantics needed by a halo-reusing 3x3 convolution:
// can one K-major SWIZZLE_128B tile, written by a single TMA box of R consecutive 128-byte rows, be
// read by an MMA whose 8-row groups start at an arbitrary row (start address offset by s*128 bytes,
// "matrix base offset" = s) and lie `sbo` bytes apart?  D[128,32] = A_shifted[128,64] * W[32,64]^T.
// Row r = 8g + i of the A operand is expected to be source row  g*(sbo/128) + row0 + i.
// Measurement probes, NOT part of libvda_b200.so: built on demand into scripts/probes/libvda_probes.so
// (scripts/probes/probes.py) and linked against the product library for its TMA-descriptor helper.
#include "../../vda_b200/csrc/common.cuh"
#include "../../vda_b200/csrc/api_internal.h"

extern "C" {
/* D f32 [128,32] = A[128,64] * W[32,64]^T where A's row 8g+i is row g*(sbo/128) + row0 + i of the bf16 source
 * G [320,64], read through a tap-shifted tcgen05 descriptor. */
int vda_debug_umma_shift(const void* G, const void* W, float* D, int row0, int sbo, int base_offset, void* stream);
/* `ctas` x `threads` threads each issue iters*8 mma.sync.m16n8k16 bf16 (no memory traffic). */
int vda_debug_mma_rate(float* out, int ctas, int threads, int iters, void* stream);
/* Pipe-rate probes for the attention kernel's softmax warps: every warp of `ctas` x `threads` runs `iters` rounds of
 * what = 0: four tcgen05.ld.32x32b.x32 (its 32 lanes x 128 fp32 columns = 16 KB) + wait::ld;
 * what = 1: 32 independent ex2.approx.ftz.f32;  what = 2: 32 independent cvt.rn.bf16x2.f32;
 * what = 3: 16 ex2 + 8 cvt + 8 fma.f32x2 interleaved (the softmax mix).
 * clk[cta*warps + warp] = clock64 ticks that warp spent in the loop. */
int vda_debug_pipe_rate(long long* clk, int what, int ctas, int threads, int iters, void* stream);
}

namespace vda {

constexpr int kProbeRows = 320;  // rows of the source tile held in smem (two TMA boxes of 160)

__global__ void __launch_bounds__(128) umma_probe_kernel(const __grid_constant__ CUtensorMap tmG,
                                                         const __grid_constant__ CUtensorMap tmW, float* __restrict__ D,
                                                         int row0, int sbo, int base_offset) {
  extern __shared__ __align__(1024) uint8_t psm[];
  uint8_t* sA = psm;                          // kProbeRows x 128 B
  uint8_t* sB = psm + kProbeRows * 128;       // 32 x 128 B (1024-aligned: 320*128 = 40 KB)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + 32 * 128);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, 32);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bars[0], kProbeRows * 128 + 32 * 128);
    tma_load_2d(sA, &tmG, &bars[0], 0, 0);
    tma_load_2d(sA + 160 * 128, &tmG, &bars[0], 0, 160);
    tma_load_2d(sB, &tmW, &bars[0], 0, 0);
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    const uint32_t a_addr = smem_u32(sA) + static_cast<uint32_t>(row0) * 128;
    const uint32_t lo_a = ((a_addr >> 4) & 0x3FFFu) | (1u << 16);
    const uint32_t hi_a = (static_cast<uint32_t>(sbo >> 4) & 0x3FFFu) | (1u << 14) | (static_cast<uint32_t>(base_offset & 7) << 17) | (2u << 29);
    const uint64_t adesc = (static_cast<uint64_t>(hi_a) << 32) | lo_a;
    const uint64_t bdesc = umma_desc_k_sw128(smem_u32(sB));
    constexpr uint32_t idesc = umma_idesc_f16(128, 32, true);
#pragma unroll
    for (int k = 0; k < 4; ++k) umma_f16_ss(tmem, adesc + 2 * k, bdesc + 2 * k, idesc, k != 0);
    umma_commit(&bars[1]);
  }
  __syncthreads();
  mbar_wait(&bars[1], 0);
  tc_fence_after();
  uint32_t v[32];
  tmem_ld32(tmem + (static_cast<uint32_t>(warp * 32) << 16), v);
  tmem_ld_wait();
  const int r = warp * 32 + lane;
#pragma unroll
  for (int c = 0; c < 32; ++c) D[r * 32 + c] = __uint_as_float(v[c]);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 32);
  }
}

// Issue-rate probe of the legacy warp-level tensor path: every warp runs `iters` rounds of 8 independent
// mma.sync.m16n8k16 bf16 chains (no memory traffic).  Sets the ceiling for the split-bf16 attention kernels.
__global__ void __launch_bounds__(512) mma_rate_kernel(float* out, int iters) {
  float c[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
  uint32_t a[4] = {threadIdx.x, threadIdx.x * 3u, threadIdx.x * 5u, threadIdx.x * 7u};
  uint32_t b0 = threadIdx.x * 11u, b1 = threadIdx.x * 13u;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                   : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                   : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  if (acc == 123.456f) out[0] = acc;
}


template <int what>
__global__ void __launch_bounds__(512) pipe_rate_kernel(long long* clk, int iters) {
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    tmem_alloc(&tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t t0 = tmem_slot + (static_cast<uint32_t>((warp & 3) * 32) << 16) + ((warp >> 2) & 3) * 128;
  uint32_t acc = lane;
  float f[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) f[i] = -0.001f * (lane + i + 1);
  __syncthreads();
  const long long c0 = clock64();
  if (what == 0) {
    for (int it = 0; it < iters; ++it) {
      uint32_t a[32], b[32], c[32], d[32];
      tmem_ld32(t0, a);
      tmem_ld32(t0 + 32, b);
      tmem_ld32(t0 + 64, c);
      tmem_ld32(t0 + 96, d);
      tmem_ld_wait();
      acc ^= a[0] ^ b[7] ^ c[13] ^ d[31];
    }
  } else if (what == 1) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 32; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
    }
  } else if (what == 2) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        uint32_t h0, h1;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h0) : "f"(f[i]), "f"(f[i + 1]));
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h1) : "f"(f[i + 1]), "f"(f[i]));
        acc ^= h0 + h1;
      }
    }
  } else if (what == 4 || what == 5 || what == 8) {
    float m[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) m[i] = f[i];
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        if (what == 4) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(m[i & 7]) : "f"(f[i]), "f"(f[(i + 5) & 31]));
        if (what == 5) asm volatile("max.f32 %0, %0, %1;" : "+f"(m[i & 7]) : "f"(f[i]));
        if (what == 8) {
          if (i & 1) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(m[i & 7]) : "f"(f[i]), "f"(f[(i + 5) & 31]));
          else asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[i]));
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc ^= __float_as_uint(m[i]);
  } else if (what == 6 || what == 7) {
    uint64_t g[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) asm("mov.b64 %0, {%1, %2};" : "=l"(g[i]) : "f"(f[2 * i]), "f"(f[2 * i + 1]));
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        if (what == 6) asm volatile("fma.rn.f32x2 %0, %0, %0, %0;" : "+l"(g[i & 15]));
        else asm volatile("add.f32x2 %0, %0, %0;" : "+l"(g[i & 15]));
      }
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) acc ^= static_cast<uint32_t>(g[i]);
  } else {
    uint64_t g[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) asm("mov.b64 %0, {%1, %2};" : "=l"(g[i]) : "f"(f[2 * i]), "f"(f[2 * i + 1]));
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[2 * i]));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[2 * i + 1]));
        uint32_t h0;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(h0) : "f"(f[16 + i]), "f"(f[17 + i]));
        asm volatile("fma.rn.f32x2 %0, %0, %0, %0;" : "+l"(g[i]));
        acc ^= h0;
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc ^= static_cast<uint32_t>(g[i]);
  }
  const long long c1 = clock64();
#pragma unroll
  for (int i = 0; i < 32; ++i) acc ^= __float_as_uint(f[i]);
  if (lane == 0) clk[blockIdx.x * (blockDim.x >> 5) + warp] = (c1 - c0) | (acc == 0x12345u ? 1 : 0);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_slot, 512);
  }
}

}  // namespace vda

using namespace vda;

extern "C" int vda_debug_pipe_rate(long long* clk, int what, int ctas, int threads, int iters, void* stream) {
  if (!clk || ctas <= 0 || threads <= 0 || threads > 512 || threads % 32) return set_error(-1, "vda_debug_pipe_rate: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (what) {
#define VDA_PR(W) case W: pipe_rate_kernel<W><<<ctas, threads, 0, st>>>(clk, iters); break;
    VDA_PR(0) VDA_PR(1) VDA_PR(2) VDA_PR(3) VDA_PR(4) VDA_PR(5) VDA_PR(6) VDA_PR(7) VDA_PR(8)
#undef VDA_PR
    default: return set_error(-1, "vda_debug_pipe_rate: what = 0..8");
  }
  return check_cuda(cudaGetLastError(), "pipe rate launch");
}

extern "C" int vda_debug_mma_rate(float* out, int ctas, int threads, int iters, void* stream) {
  if (!out || ctas <= 0 || threads <= 0 || threads > 512 || threads % 32) return set_error(-1, "vda_debug_mma_rate: bad arguments");
  mma_rate_kernel<<<ctas, threads, 0, static_cast<cudaStream_t>(stream)>>>(out, iters);
  return check_cuda(cudaGetLastError(), "mma rate launch");
}

extern "C" int vda_debug_umma_shift(const void* G, const void* W, float* D, int row0, int sbo, int base_offset, void* stream) {
  if (!G || !W || !D || row0 < 0 || sbo % 16) return set_error(-1, "vda_debug_umma_shift: bad arguments");
  CUtensorMap tmG, tmW;
  int rc;
  {
    const uint64_t dims[2] = {64, (uint64_t)kProbeRows};
    const uint64_t str[1] = {128};
    const uint32_t box[2] = {64, 160};
    if ((rc = encode_bf16_map(&tmG, G, 2, dims, str, box))) return rc;
  }
  {
    const uint64_t dims[2] = {64, 32};
    const uint64_t str[1] = {128};
    const uint32_t box[2] = {64, 32};
    if ((rc = encode_bf16_map(&tmW, W, 2, dims, str, box))) return rc;
  }
  const int smem = kProbeRows * 128 + 32 * 128 + 64 + 1024;
  rc = check_cuda(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem), "probe smem");
  if (rc) return rc;
  umma_probe_kernel<<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(tmG, tmW, D, row0, sbo, base_offset);
  return check_cuda(cudaGetLastError(), "umma probe launch");
}

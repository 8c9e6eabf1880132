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

}  // namespace vda

using namespace vda;

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

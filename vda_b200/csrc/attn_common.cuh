// Helpers shared by the spatial-attention kernels (attn.cu: one CTA per query tile; attn3.cu: persistent ping-pong).
#pragma once
#include "common.cuh"

namespace vda {

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 2^x on the FMA pipe (Cody-Waite split + degree-3 polynomial, rel. error 1.1e-4, invisible after the
// bf16 rounding of P): the MUFU unit does 16 ex2/clk/SM and bounds head_dim-64 attention, so a fixed
// fraction of the exponentials is evaluated here instead to balance the XU and FMA pipes.
__device__ __forceinline__ float poly_exp2(float x) {
  x = fmaxf(x, -126.0f);
  const float fl = x + 12582912.0f;           // 1.5 * 2^23: integer part lands in the low mantissa bits
  const float f = x - (fl - 12582912.0f);     // f in [-0.5, 0.5]
  float r = fmaf(0.05459282f, f, 0.24221784f);
  r = fmaf(r, f, 0.6933686f);
  r = fmaf(r, f, 1.0f);
  return __int_as_float(__float_as_int(r) + (__float_as_int(fl) << 23));
}

// Packed fp32x2 arithmetic (FFMA2 / FADD2) and the 3-input max (FMNMX3) of sm_100: the softmax warps are
// issue-slot bound next to the MUFU limit, and these halve the FMA-pipe instruction count per score.
__device__ __forceinline__ uint64_t pack2(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

// poly_exp2 on a pair of scores with packed arithmetic (10 FMA-pipe instructions per pair).
__device__ __forceinline__ void poly_exp2_x2(uint64_t x, float& p0, float& p1) {
  float x0, x1;
  unpack2(x, x0, x1);
  x = pack2(fmaxf(x0, -126.0f), fmaxf(x1, -126.0f));
  const uint64_t magic = pack2(12582912.0f, 12582912.0f), nmagic = pack2(-12582912.0f, -12582912.0f);
  const uint64_t fl = fadd2(x, magic);
  const uint64_t f = ffma2(fadd2(fl, nmagic), pack2(-1.0f, -1.0f), x);
  uint64_t r = ffma2(pack2(0.05459282f, 0.05459282f), f, pack2(0.24221784f, 0.24221784f));
  r = ffma2(r, f, pack2(0.6933686f, 0.6933686f));
  r = ffma2(r, f, pack2(1.0f, 1.0f));
  float r0, r1, f0, f1;
  unpack2(r, r0, r1);
  unpack2(fl, f0, f1);
  p0 = __int_as_float(__float_as_int(r0) + (__float_as_int(f0) << 23));
  p1 = __int_as_float(__float_as_int(r1) + (__float_as_int(f1) << 23));
}

constexpr int kAttThreads = 192;
constexpr int kQBytes = 128 * 64 * 2;        // 16 KB
constexpr int kKBytes = 128 * 64 * 2;        // 16 KB
constexpr int kVBytes = 64 * 128 * 2;        // 16 KB (two 64x64 sub-tiles)
constexpr int kAttTmemCols = 256;            // S: cols [0,128), P: [128,192), O: [192,256)

struct alignas(64) AttnParams {
  CUtensorMap tmQK;  // 3D (2D, S, frames), box (64, 128, 1)
  CUtensorMap tmV;   // 3D (s_pad, 64, frames*heads), box (64, 64, 1)
  bf16* out;
  int S, D, heads, frames;
  int q_tiles, kv_tiles;
  float scale_log2;  // head_dim^-0.5 * log2(e)
  int* flags;        // [frames*heads*q_tiles] overflow flags of the optimistic kernel / work list of the fix-up pass
  int ipc;           // v3 chained mode: consecutive work items per CTA
  long long* prof;   // PROF build only: [ctas*4][8] clock sums per softmax warp (wait S, ld, max, exp, wait PV, st, total)
};
__device__ __forceinline__ long long clk_after(float dep) {
  long long t;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) : "f"(dep) : "memory");
  return t;
}


}  // namespace vda

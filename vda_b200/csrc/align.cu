// Window alignment kernels (video_depth.py:120-160, utils/util.py:40-74): least-squares
// scale/shift between overlapping key frames as a GPU reduction (FP64 accumulation), 2x2 solve on
// the device (no host sync in the sequential scan), and fused affine + clamp(>=0) + cross-fade.
#include <string.h>

#include "common.cuh"
#include "api_internal.h"
#include "../../include/vda_b200.h"

namespace vda {

// sums[0]=sum p*p, [1]=sum p, [2]=count, [3]=sum p*t, [4]=sum t   (mask == 1 everywhere)
__global__ void lsq_accum_kernel(const float* __restrict__ pred, const float* __restrict__ target, long long n,
                                 double* __restrict__ sums) {
  double a00 = 0, a01 = 0, b0 = 0, b1 = 0;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float p = pred[i], t = target[i];
    a00 += static_cast<double>(p) * p;
    a01 += p;
    b0 += static_cast<double>(p) * t;
    b1 += t;
  }
  __shared__ double sh[4][8];
  double vals[4] = {a00, a01, b0, b1};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    double v = vals[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[k][threadIdx.x >> 5] = v;
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double v = 0;
    for (int w = 0; w < static_cast<int>(blockDim.x >> 5); ++w) v += sh[threadIdx.x][w];
    const int slot = threadIdx.x == 0 ? 0 : (threadIdx.x == 1 ? 1 : (threadIdx.x == 2 ? 3 : 4));
    atomicAdd(&sums[slot], v);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&sums[2], static_cast<double>(n));
}

// utils/util.py:53-60: det = a00*a11 - a01^2; det != 0 guard kept; metric models skip (1, 0).
__global__ void lsq_solve_kernel(const double* __restrict__ sums, float* __restrict__ scale_shift, int metric) {
  float s = 1.f, t = 0.f;
  if (!metric) {
    const double a00 = sums[0], a01 = sums[1], a11 = sums[2], b0 = sums[3], b1 = sums[4];
    const double det = a00 * a11 - a01 * a01;
    if (det != 0) {
      s = static_cast<float>((a11 * b0 - a01 * b1) / det);
      t = static_cast<float>((-a01 * b0 + a00 * b1) / det);
    }
  }
  scale_shift[0] = s;
  scale_shift[1] = t;
}

// dst = max(src*s + t, 0)                                        (video_depth.py:143-144,149-151)
__global__ void affine_clamp_kernel(const float* __restrict__ src, float* __restrict__ dst, long long n,
                                    const float* __restrict__ scale_shift) {
  const float s = scale_shift[0], t = scale_shift[1];
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = fmaxf(src[i] * s + t, 0.f);
}

// pre[f] = pre[f]*(1-w_f) + max(post[f]*s+t, 0)*w_f for the 8 overlapping frames (utils/util.py:65-74)
struct BlendW {
  float w[8], omw[8];
};
__global__ void affine_clamp_blend_kernel(float* __restrict__ pre, const float* __restrict__ post, long long per_frame,
                                          int frames, const float* __restrict__ scale_shift, BlendW bw) {
  const float s = scale_shift[0], t = scale_shift[1];
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= per_frame * frames) return;
  const int f = static_cast<int>(i / per_frame);
  const float a = fmaxf(post[i] * s + t, 0.f);
  pre[i] = pre[i] * bw.omw[f] + a * bw.w[f];
}

}  // namespace vda

using namespace vda;

extern "C" {

int vda_lsq_accum(const float* pred, const float* target, long long n, double* sums5, void* stream) {
  if (!pred || !target || !sums5 || n <= 0) return set_error(-1, "vda_lsq_accum: bad arguments");
  int blocks = static_cast<int>((n + 256 * 8 - 1) / (256 * 8));
  if (blocks > 1184) blocks = 1184;
  lsq_accum_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(pred, target, n, sums5);
  return check_cuda(cudaGetLastError(), "lsq_accum launch");
}

int vda_lsq_solve(const double* sums5, float* scale_shift, int metric, void* stream) {
  if (!sums5 || !scale_shift) return set_error(-1, "vda_lsq_solve: bad arguments");
  lsq_solve_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(sums5, scale_shift, metric);
  return check_cuda(cudaGetLastError(), "lsq_solve launch");
}

int vda_affine_clamp(const float* src, float* dst, long long n, const float* scale_shift, void* stream) {
  if (!src || !dst || !scale_shift || n <= 0) return set_error(-1, "vda_affine_clamp: bad arguments");
  affine_clamp_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(src, dst, n,
                                                                                                           scale_shift);
  return check_cuda(cudaGetLastError(), "affine_clamp launch");
}

int vda_affine_clamp_blend(float* pre, const float* post, long long per_frame, int frames, const float* scale_shift,
                           void* stream) {
  if (!pre || !post || !scale_shift || frames < 2 || frames > 8) return set_error(-1, "vda_affine_clamp_blend: bad arguments");
  BlendW bw;
  const double step = 1.0 / (frames - 1);
  for (int i = 0; i < frames; ++i) {
    const double w = (i == 0) ? 0.0 : (i == frames - 1 ? 1.0 : i * step);
    bw.w[i] = static_cast<float>(w);
    bw.omw[i] = static_cast<float>(1.0 - w);
  }
  const long long n = per_frame * frames;
  affine_clamp_blend_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      pre, post, per_frame, frames, scale_shift, bw);
  return check_cuda(cudaGetLastError(), "affine_clamp_blend launch");
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Depth visualisation (utils/dc_utils.py:72-86: global min/max over the whole clip, 8-bit
// quantisation, "inferno" look-up or grey).  Two bandwidth kernels: a min/max reduction (depth >= 0,
// so the IEEE bit patterns order like unsigned integers) and the per-pixel LUT.
namespace vda {

__global__ void minmax_init_kernel(unsigned int* mm) { mm[0] = 0x7f800000u, mm[1] = 0u; }  // +inf, 0

__global__ void minmax_f32_kernel(const float* __restrict__ x, long long n, unsigned int* __restrict__ mm) {
  float lo = INFINITY, hi = 0.f;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float v = fmaxf(x[i], 0.f);
    lo = fminf(lo, v), hi = fmaxf(hi, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(mm, __float_as_uint(lo));
    atomicMax(mm + 1, __float_as_uint(hi));
  }
}

// out[i] = lut[q] (3 bytes) or q (grey), q = uint8((d - min) / (max - min) * 255) with numpy's truncating cast
__global__ void depth_colorize_kernel(const float* __restrict__ d, long long n, const unsigned int* __restrict__ mm,
                                      const uint8_t* __restrict__ lut, uint8_t* __restrict__ out, int grey) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float lo = __uint_as_float(mm[0]), hi = __uint_as_float(mm[1]);
  const float q = (d[i] - lo) / (hi - lo) * 255.f;
  const int k = min(max(static_cast<int>(q), 0), 255);
  if (grey) {
    out[i] = static_cast<uint8_t>(k);
  } else {
    out[3 * i] = lut[3 * k], out[3 * i + 1] = lut[3 * k + 1], out[3 * i + 2] = lut[3 * k + 2];
  }
}

}  // namespace vda

extern "C" int vda_minmax_f32(const float* x, long long n, void* minmax2, void* stream) {
  if (!x || !minmax2 || n <= 0) return set_error(-1, "vda_minmax_f32: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  vda::minmax_init_kernel<<<1, 1, 0, st>>>(static_cast<unsigned int*>(minmax2));
  const int blocks = static_cast<int>(n / 1024 < 1184 ? (n + 1023) / 1024 : 1184);
  vda::minmax_f32_kernel<<<blocks, 256, 0, st>>>(x, n, static_cast<unsigned int*>(minmax2));
  return vda::check_cuda(cudaGetLastError(), "minmax launch");
}

extern "C" int vda_depth_colorize(const float* depth, long long n, const void* minmax2, const void* lut_rgb, void* out,
                                  int grey, void* stream) {
  if (!depth || !minmax2 || !out || n <= 0 || (!grey && !lut_rgb)) return set_error(-1, "vda_depth_colorize: bad arguments");
  vda::depth_colorize_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      depth, n, static_cast<const unsigned int*>(minmax2), static_cast<const uint8_t*>(lut_rgb), static_cast<uint8_t*>(out), grey);
  return vda::check_cuda(cudaGetLastError(), "depth_colorize launch");
}

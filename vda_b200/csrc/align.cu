// Window alignment kernels (video_depth.py:120-160, utils/util.py:40-74): least-squares
// scale/shift between overlapping key frames as a GPU reduction (FP64 accumulation), 2x2 solve on
// the device (no host sync in the sequential scan), and fused affine + clamp(>=0) + cross-fade.
#include <string.h>

#include "common.cuh"
#include "api_internal.h"
#include "../../include/vda_b200.h"

namespace vda {

// sums[0]=sum m*p*p, [1]=sum m*p, [2]=sum m, [3]=sum m*p*t, [4]=sum m*t; mask == nullptr means m == 1 everywhere
// (the inference path, video_depth.py:138-140), otherwise m is an fp32 weight per element (utils/util.py:43-51).
template <bool MASKED>
__global__ void lsq_accum_kernel(const float* __restrict__ pred, const float* __restrict__ target,
                                 const float* __restrict__ mask, long long n, double* __restrict__ sums) {
  double a00 = 0, a01 = 0, b0 = 0, b1 = 0, a11 = 0;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float p = pred[i], t = target[i];
    if (MASKED) {
      const double m = mask[i];
      a00 += m * p * p;
      a01 += m * p;
      b0 += m * p * t;
      b1 += m * t;
      a11 += m;
    } else {
      a00 += static_cast<double>(p) * p;
      a01 += p;
      b0 += static_cast<double>(p) * t;
      b1 += t;
    }
  }
  __shared__ double sh[5][8];
  double vals[5] = {a00, a01, b0, b1, a11};
#pragma unroll
  for (int k = 0; k < (MASKED ? 5 : 4); ++k) {
    double v = vals[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[k][threadIdx.x >> 5] = v;
  }
  __syncthreads();
  if (threadIdx.x < (MASKED ? 5 : 4)) {
    double v = 0;
    for (int w = 0; w < static_cast<int>(blockDim.x >> 5); ++w) v += sh[threadIdx.x][w];
    const int slot = threadIdx.x == 0 ? 0 : (threadIdx.x == 1 ? 1 : (threadIdx.x == 2 ? 3 : (threadIdx.x == 3 ? 4 : 2)));
    atomicAdd(&sums[slot], v);
  }
  if (!MASKED && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&sums[2], static_cast<double>(n));
}

// utils/util.py:53-60: det = a00*a11 - a01^2; det != 0 guard kept; metric models skip (1, 0).
// flags bit 0: metric model (scale 1, shift 0); bit 1: zero the five sums afterwards, ready for the next window.
__global__ void lsq_solve_kernel(double* __restrict__ sums, float* __restrict__ scale_shift, int flags) {
  float s = 1.f, t = 0.f;
  if (!(flags & 1)) {
    const double a00 = sums[0], a01 = sums[1], a11 = sums[2], b0 = sums[3], b1 = sums[4];
    const double det = a00 * a11 - a01 * a01;
    if (det != 0) {
      s = static_cast<float>((a11 * b0 - a01 * b1) / det);
      t = static_cast<float>((-a01 * b0 + a00 * b1) / det);
    }
  }
  scale_shift[0] = s;
  scale_shift[1] = t;
  if (flags & 2) {
#pragma unroll
    for (int i = 0; i < 5; ++i) sums[i] = 0.0;
  }
}



// dst = max(src*s + t, 0)                                        (video_depth.py:143-144,149-151)
__global__ void affine_clamp_kernel(const float* __restrict__ src, float* __restrict__ dst, long long n,
                                    const float* __restrict__ scale_shift) {
  const float s = scale_shift[0], t = scale_shift[1];
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = fmaxf(src[i] * s + t, 0.f);
}

// pre[f] = pre[f]*(1-w_f) + max(post[f]*s+t, 0)*w_f for the 8 overlapping frames (utils/util.py:65-74)
// blend_frames_kernel: out[f] = pre[f]*(1-w_f) + post[f]*w_f, each product and the sum rounded to fp32 like
// numpy's (utils/util.py:65-74), without the affine / clamp.
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

__global__ void blend_frames_kernel(const float* __restrict__ pre, const float* __restrict__ post, float* __restrict__ out,
                                    long long per_frame, int frames, BlendW bw) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= per_frame * frames) return;
  const int f = static_cast<int>(i / per_frame);
  out[i] = __fadd_rn(__fmul_rn(pre[i], bw.omw[f]), __fmul_rn(post[i], bw.w[f]));
}

static BlendW blend_weights(int frames) {
  BlendW bw;
  const double step = 1.0 / (frames - 1);
  for (int i = 0; i < frames; ++i) {
    const double w = (i == 0) ? 0.0 : (i == frames - 1 ? 1.0 : i * step);
    bw.w[i] = static_cast<float>(w);
    bw.omw[i] = static_cast<float>(1.0 - w);
  }
  return bw;
}

}  // namespace vda

using namespace vda;

extern "C" {

int vda_lsq_accum(const float* pred, const float* target, long long n, double* sums5, void* stream) {
  return vda_lsq_accum_masked(pred, target, nullptr, n, sums5, stream);
}

int vda_lsq_accum_masked(const float* pred, const float* target, const float* mask, long long n, double* sums5, void* stream) {
  if (!pred || !target || !sums5 || n <= 0) return set_error(-1, "vda_lsq_accum: bad arguments");
  int blocks = static_cast<int>((n + 256 * 8 - 1) / (256 * 8));
  if (blocks > 1184) blocks = 1184;
  if (mask)
    lsq_accum_kernel<true><<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(pred, target, mask, n, sums5);
  else
    lsq_accum_kernel<false><<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(pred, target, nullptr, n, sums5);
  return check_cuda(cudaGetLastError(), "lsq_accum launch");
}

int vda_lsq_solve(double* sums5, float* scale_shift, int flags, void* stream) {
  if (!sums5 || !scale_shift) return set_error(-1, "vda_lsq_solve: bad arguments");
  lsq_solve_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(sums5, scale_shift, flags);
  return check_cuda(cudaGetLastError(), "lsq_solve launch");
}

int vda_blend_frames(const float* pre, const float* post, float* out, long long per_frame, int frames, void* stream) {
  if (!pre || !post || !out || frames < 2 || frames > 8 || per_frame <= 0) return set_error(-1, "vda_blend_frames: bad arguments");
  const long long n = per_frame * frames;
  blend_frames_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      pre, post, out, per_frame, frames, blend_weights(frames));
  return check_cuda(cudaGetLastError(), "blend_frames launch");
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
  const BlendW bw = blend_weights(frames);
  const long long n = per_frame * frames;
  affine_clamp_blend_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      pre, post, per_frame, frames, scale_shift, bw);
  return check_cuda(cudaGetLastError(), "affine_clamp_blend launch");
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Depth visualisation and normalisation used by the reference's callers:
//   * utils/dc_utils.py:72-86 -- clip-global min/max, 8-bit quantisation, colour look-up or grey;
//   * external_engine/engine.py:1099-1141 (`normalize_depth`) -- per-frame min/max -> [near, far] remap,
//     inverted mapping when far <= near, `invert_depth`, constant-depth frames set to `near`.
// Bandwidth kernels: a (batched) min/max reduction and a per-pixel map.  Every float operation of the
// reference is float32 with one rounding each, so the kernels use the non-contracting intrinsics
// (__fsub_rn / __fdiv_rn / __fmul_rn / __fadd_rn): the bytes / floats are identical to numpy's.
namespace vda {

// order-preserving map float -> unsigned (any sign), so atomicMin / atomicMax work on the keys
__device__ __forceinline__ unsigned int f2key(float f) {
  const unsigned int b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key2f(unsigned int k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__global__ void minmax_init_kernel(unsigned int* mm, int batches) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < batches) mm[2 * i] = 0xffffffffu, mm[2 * i + 1] = 0u;
}

// mm[2b], mm[2b+1] = keys of min / max over x[b*per : (b+1)*per], b = blockIdx.y
__global__ void minmax_f32_kernel(const float* __restrict__ x, long long per, unsigned int* __restrict__ mm) {
  const float* xb = x + static_cast<long long>(blockIdx.y) * per;
  float lo = INFINITY, hi = -INFINITY;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < per;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float v = xb[i];
    lo = fminf(lo, v), hi = fmaxf(hi, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(mm + 2 * blockIdx.y, f2key(lo));
    atomicMax(mm + 2 * blockIdx.y + 1, f2key(hi));
  }
}

// keys -> the two floats, in place (one thread per batch), so callers can read {min, max} as fp32
__global__ void minmax_decode_kernel(unsigned int* mm, int batches) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 2 * batches) mm[i] = __float_as_uint(key2f(mm[i]));
}

// out[i] = lut[q] (3 bytes) or q (grey), q = ((d - min) / (max - min) * 255).astype(uint8): three fp32 roundings,
// truncating cast (dc_utils.py:78)
__global__ void depth_colorize_kernel(const float* __restrict__ d, long long n, const float* __restrict__ mm,
                                      const uint8_t* __restrict__ lut, uint8_t* __restrict__ out, int grey) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float lo = mm[0], hi = mm[1];
  const float q = __fmul_rn(__fdiv_rn(__fsub_rn(d[i], lo), __fsub_rn(hi, lo)), 255.f);
  const int k = static_cast<int>(q) & 255;   // numpy's float -> uint8 cast: truncate, then wrap modulo 256
  if (grey) {
    out[i] = static_cast<uint8_t>(k);
  } else {
    out[3 * i] = lut[3 * k], out[3 * i + 1] = lut[3 * k + 1], out[3 * i + 2] = lut[3 * k + 2];
  }
}

// engine.py:1110-1129 per frame b (mm[2b], mm[2b+1] = that frame's min / max)
__global__ void depth_normalize_kernel(const float* __restrict__ d, float* __restrict__ out, long long per,
                                       const float* __restrict__ mm, float near_v, float far_v, float base, float span,
                                       int invert) {
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= per) return;
  const long long at = static_cast<long long>(blockIdx.y) * per + i;
  const float lo = mm[2 * blockIdx.y], hi = mm[2 * blockIdx.y + 1];
  float v;
  if (hi > lo) {
    const float t = __fdiv_rn(__fsub_rn(d[at], lo), __fsub_rn(hi, lo));
    v = __fadd_rn(base, __fmul_rn(span, t));   // near + (far-near)*t, or far + (near-far)*t when far <= near
  } else {
    v = near_v;                                // constant depth
  }
  if (invert) v = __fsub_rn(far_v, __fsub_rn(v, near_v));
  out[at] = v;
}

static int launch_minmax(const float* x, long long per, int batches, void* mm, cudaStream_t st) {
  minmax_init_kernel<<<(batches + 255) / 256, 256, 0, st>>>(static_cast<unsigned int*>(mm), batches);
  int bx = static_cast<int>(per / 1024 < 1184 ? (per + 1023) / 1024 : 1184);
  if (batches > 1 && bx > 148) bx = 148;
  minmax_f32_kernel<<<dim3(bx, batches), 256, 0, st>>>(x, per, static_cast<unsigned int*>(mm));
  minmax_decode_kernel<<<(2 * batches + 255) / 256, 256, 0, st>>>(static_cast<unsigned int*>(mm), batches);
  return check_cuda(cudaGetLastError(), "minmax launch");
}

}  // namespace vda

extern "C" int vda_minmax_f32(const float* x, long long n, void* minmax2, void* stream) {
  if (!x || !minmax2 || n <= 0) return vda::set_error(-1, "vda_minmax_f32: bad arguments");
  return vda::launch_minmax(x, n, 1, minmax2, static_cast<cudaStream_t>(stream));
}

extern "C" int vda_depth_colorize(const float* depth, long long n, const void* minmax2, const void* lut_rgb, void* out,
                                  int grey, void* stream) {
  if (!depth || !minmax2 || !out || n <= 0 || (!grey && !lut_rgb)) return vda::set_error(-1, "vda_depth_colorize: bad arguments");
  vda::depth_colorize_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      depth, n, static_cast<const float*>(minmax2), static_cast<const uint8_t*>(lut_rgb), static_cast<uint8_t*>(out), grey);
  return vda::check_cuda(cudaGetLastError(), "depth_colorize launch");
}

extern "C" int vda_depth_normalize(const float* depth, float* out, long long per_frame, int frames, double near_value,
                                   double far_value, int invert, void* minmax_ws, void* stream) {
  if (!depth || !out || !minmax_ws || per_frame <= 0 || frames <= 0 || frames > 65535)
    return vda::set_error(-1, "vda_depth_normalize: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = vda::launch_minmax(depth, per_frame, frames, minmax_ws, st);
  if (rc) return rc;
  // python-float scalars meet float32 arrays: the scalar expressions are evaluated in double, then cast once
  const bool fwd = far_value > near_value;
  const float base = static_cast<float>(fwd ? near_value : far_value);
  const float span = static_cast<float>(fwd ? far_value - near_value : near_value - far_value);
  vda::depth_normalize_kernel<<<dim3(static_cast<unsigned>((per_frame + 255) / 256), frames), 256, 0, st>>>(
      depth, out, per_frame, static_cast<const float*>(minmax_ws), static_cast<float>(near_value),
      static_cast<float>(far_value), base, span, invert);
  return vda::check_cuda(cudaGetLastError(), "depth_normalize launch");
}

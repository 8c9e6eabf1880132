// Frame pre-processing on the GPU (util/transform.py:51-157 as composed in video_depth.py:77-89):
//   uint8 RGB HWC frame -> float32 /255 -> cv2.resize(INTER_CUBIC) -> (x - mean) / std in float64 ->
//   float32 CHW network input.
// The resize restates OpenCV's float32 cubic path (cv2 is the third-party arithmetic behind
// transform.py:113): source coordinate fx = (dx + 0.5) * (src/dst) - 0.5, 4 taps at
// floor(fx) - 1 .. + 2 with Keys weights for A = -0.75 evaluated in float32, tap indices clamped to
// the image (replicate border), no antialiasing; horizontal pass first, then vertical, as cv2 does.
// One thread per output pixel (all 3 channels): HBM-bound, 48 source bytes (L1/L2 hits across
// neighbours) and 12 output bytes per pixel.
#include "common.cuh"
#include "api_internal.h"
#include "../../include/vda_b200.h"

namespace vda {

__device__ __forceinline__ void cubic_coeffs(float x, float (&c)[4]) {
  const float A = -0.75f;
  c[0] = ((A * (x + 1.f) - 5.f * A) * (x + 1.f) + 8.f * A) * (x + 1.f) - 4.f * A;
  c[1] = ((A + 2.f) * x - (A + 3.f)) * x * x + 1.f;
  c[2] = ((A + 2.f) * (1.f - x) - (A + 3.f)) * (1.f - x) * (1.f - x) + 1.f;
  c[3] = 1.f - c[0] - c[1] - c[2];
}

__global__ void __launch_bounds__(256) preprocess_u8_kernel(const uint8_t* __restrict__ frames, float* __restrict__ out, int H0,
                                                            int W0, int nh, int nw, double scale_y, double scale_x) {
  const int X = blockIdx.x * blockDim.x + threadIdx.x;
  const int Y = blockIdx.y, f = blockIdx.z;
  if (X >= nw) return;
  float fx = static_cast<float>((X + 0.5) * scale_x - 0.5);
  float fy = static_cast<float>((Y + 0.5) * scale_y - 0.5);
  const int sx = static_cast<int>(floorf(fx)), sy = static_cast<int>(floorf(fy));
  fx -= sx, fy -= sy;
  float cx[4], cy[4];
  cubic_coeffs(fx, cx);
  cubic_coeffs(fy, cy);
  int xs[4], ys[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    xs[k] = min(max(sx + k - 1, 0), W0 - 1);
    ys[k] = min(max(sy + k - 1, 0), H0 - 1);
  }
  const uint8_t* img = frames + static_cast<long long>(f) * H0 * W0 * 3;
  float acc[3] = {0.f, 0.f, 0.f};
  const bool identity = (nh == H0 && nw == W0);  // cv2.resize to the same size is a copy
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint8_t* row = img + static_cast<long long>(ys[k]) * W0 * 3;
    float r[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint8_t* px = row + xs[j] * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) r[c] = __fadd_rn(r[c], __fmul_rn(static_cast<float>(px[c]) / 255.0f, cx[j]));
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) acc[c] = __fadd_rn(acc[c], __fmul_rn(r[c], cy[k]));
  }
  if (identity) {
    const uint8_t* px = img + (static_cast<long long>(Y) * W0 + X) * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c) acc[c] = static_cast<float>(px[c]) / 255.0f;
  }
  const double mean[3] = {0.485, 0.456, 0.406}, stdv[3] = {0.229, 0.224, 0.225};
  float* o = out + (static_cast<long long>(f) * 3 * nh + Y) * nw + X;
#pragma unroll
  for (int c = 0; c < 3; ++c)
    o[static_cast<long long>(c) * nh * nw] = static_cast<float>((static_cast<double>(acc[c]) - mean[c]) / stdv[c]);
}

}  // namespace vda

using namespace vda;

extern "C" int vda_preprocess_u8(const void* frames, int n, int H0, int W0, float* out, int nh, int nw, void* stream) {
  if (!frames || !out || n <= 0 || H0 <= 0 || W0 <= 0 || nh <= 0 || nw <= 0 || nh > 65535 || n > 65535)
    return set_error(-1, "vda_preprocess_u8: bad arguments");
  const dim3 grid((nw + 255) / 256, nh, n);
  preprocess_u8_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const uint8_t*>(frames), out, H0, W0, nh, nw, static_cast<double>(H0) / nh, static_cast<double>(W0) / nw);
  return check_cuda(cudaGetLastError(), "preprocess_u8 launch");
}

// Temporal attention core of the motion modules (motion_module/motion_module.py:242-321,
// motion_module/attention.py:187-207): for every spatial token and each of the 8 heads,
// softmax(q k^T / sqrt(hd)) v over the T <= 32 frames of the window.  The q/k/v projections and
// to_out are plain GEMMs (gemm.cu); the "(b f) d c <-> (b d) f c" rearranges of the reference are
// pure index math here: element (t, d, c) lives at row t*hw + d.
//
// One warp per (token, head): K and V rows staged in shared memory (broadcast reads), lane t owns
// query frame t, the 32 scores and the softmax live in that lane's registers (no shuffles), the
// output row is staged back through shared memory for coalesced 16-byte stores.
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "api_internal.h"
#include "../../include/vda_b200.h"

namespace vda {

constexpr int kTWarps = 4;

struct TemporalParams {
  const bf16* q;
  const bf16* k;
  const bf16* v;
  bf16* out;
  int ldq, ldk, ldv, ldo;
  int hw, heads, hd;
  int Tq, Tk;
  float scale;
};

__device__ __forceinline__ float2 bf2(uint32_t w) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w));
}

__global__ void __launch_bounds__(kTWarps * 32) temporal_attn_kernel(const TemporalParams p) {
  extern __shared__ __align__(16) uint8_t tsm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int unit = blockIdx.x * kTWarps + warp;
  if (unit >= p.hw * p.heads) return;
  const int d = unit / p.heads, head = unit - d * p.heads;
  const int hd = p.hd, chunks = hd >> 3;  // 16-byte chunks per row
  const int q_stride = hd + 2;            // bf16 elements; odd word stride -> conflict-free per-lane rows
  const int per_warp = 32 * q_stride * 2 + 2 * 32 * hd * 2;
  uint8_t* base = tsm + warp * per_warp;
  bf16* sK = reinterpret_cast<bf16*>(base);
  bf16* sV = sK + 32 * hd;
  bf16* sQ = sV + 32 * hd;

  // cooperative 16-byte loads of the K, V (Tk rows) and Q (Tq rows) slices of this (token, head)
  for (int i = lane; i < p.Tk * chunks; i += 32) {
    const int t = i / chunks, c = i - t * chunks;
    const long long row = static_cast<long long>(t) * p.hw + d;
    const uint4 kk = __ldg(reinterpret_cast<const uint4*>(p.k + row * p.ldk + head * hd + c * 8));
    const uint4 vv = __ldg(reinterpret_cast<const uint4*>(p.v + row * p.ldv + head * hd + c * 8));
    *reinterpret_cast<uint4*>(sK + t * hd + c * 8) = kk;
    *reinterpret_cast<uint4*>(sV + t * hd + c * 8) = vv;
  }
  for (int i = lane; i < p.Tq * chunks; i += 32) {
    const int t = i / chunks, c = i - t * chunks;
    const long long row = static_cast<long long>(t) * p.hw + d;
    const uint4 qq = __ldg(reinterpret_cast<const uint4*>(p.q + row * p.ldq + head * hd + c * 8));
    uint32_t* dst = reinterpret_cast<uint32_t*>(sQ + t * q_stride + c * 8);
    dst[0] = qq.x, dst[1] = qq.y, dst[2] = qq.z, dst[3] = qq.w;
  }
  __syncwarp();

  const int t = lane;  // query frame of this lane (idle if >= Tq)
  float s[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) s[j] = 0.f;
  const uint32_t* qrow = reinterpret_cast<const uint32_t*>(sQ + (t < p.Tq ? t : 0) * q_stride);
  for (int c = 0; c < chunks; ++c) {
    const float2 q0 = bf2(qrow[c * 4]), q1 = bf2(qrow[c * 4 + 1]), q2 = bf2(qrow[c * 4 + 2]), q3 = bf2(qrow[c * 4 + 3]);
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (j < p.Tk) {
        const uint4 kk = *reinterpret_cast<const uint4*>(sK + j * hd + c * 8);
        const float2 k0 = bf2(kk.x), k1 = bf2(kk.y), k2 = bf2(kk.z), k3 = bf2(kk.w);
        s[j] += q0.x * k0.x + q0.y * k0.y + q1.x * k1.x + q1.y * k1.y + q2.x * k2.x + q2.y * k2.y + q3.x * k3.x +
                q3.y * k3.y;
      }
    }
  }
  float m = -INFINITY;
#pragma unroll
  for (int j = 0; j < 32; ++j)
    if (j < p.Tk) m = fmaxf(m, s[j] * p.scale);
  float l = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    s[j] = (j < p.Tk) ? __expf(s[j] * p.scale - m) : 0.f;
    l += s[j];
  }
  const float inv = 1.f / l;
  __syncwarp();  // every lane is done reading its q row before it is overwritten with the output
  uint32_t* orow = reinterpret_cast<uint32_t*>(sQ + (t < p.Tq ? t : 0) * q_stride);
  for (int c = 0; c < chunks; ++c) {
    float o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      if (j < p.Tk) {
        const uint4 vv = *reinterpret_cast<const uint4*>(sV + j * hd + c * 8);
        const float2 v0 = bf2(vv.x), v1 = bf2(vv.y), v2 = bf2(vv.z), v3 = bf2(vv.w);
        o[0] += s[j] * v0.x, o[1] += s[j] * v0.y, o[2] += s[j] * v1.x, o[3] += s[j] * v1.y;
        o[4] += s[j] * v2.x, o[5] += s[j] * v2.y, o[6] += s[j] * v3.x, o[7] += s[j] * v3.y;
      }
    }
    if (t < p.Tq) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(o[2 * i] * inv, o[2 * i + 1] * inv);
        orow[c * 4 + i] = *reinterpret_cast<const uint32_t*>(&h);
      }
    }
  }
  __syncwarp();
  for (int i = lane; i < p.Tq * chunks; i += 32) {
    const int tt = i / chunks, c = i - tt * chunks;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(sQ + tt * q_stride + c * 8);
    const long long row = static_cast<long long>(tt) * p.hw + d;
    *reinterpret_cast<uint4*>(p.out + row * p.ldo + head * hd + c * 8) = make_uint4(src[0], src[1], src[2], src[3]);
  }
}

// ---------------------------------------------------------------------------------------------
// Warp-MMA version of the same op (the default).  The SIMT kernel above spends ~12k FMA/convert
// instructions per (token, head) and stages K/V/Q through registers one 16-byte load at a time
// (0.75 ms for the 1369x8x(32x128) module of vitl against ~60 us of HBM time).  Here each warp
//   1. copies its Q, K, V slices (32 rows x head_dim) global -> shared with cp.async, all requests in
//      flight at once, rows padded by 16 bytes so ldmatrix is bank-conflict free,
//   2. S = Q K^T with mma.sync.m16n8k16 (bf16 in, fp32 accumulate): 2 m-tiles x 4 n-tiles,
//   3. softmax over the 32 keys in the accumulator registers (quad shuffles),
//   4. O = P V with P re-used from the accumulators as A fragments and V read with ldmatrix.trans,
//   5. O staged through the (dead) Q rows for 16-byte coalesced stores.
// The temporal axis is only 32 long, so these are warp-level tensor-core tiles, not tcgen05 ones:
// the op is HBM-bound (4 x rows x C x 2 bytes) and the MMA work is negligible.
__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void* src, int bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}

template <int HDP>  // head_dim rounded up to a multiple of 16
__global__ void __launch_bounds__(kTWarps * 32) temporal_attn_mma_kernel(const TemporalParams p) {
  pdl_sync();  // launched with programmatic stream serialization (common.cuh)
  constexpr int kStride = HDP * 2 + 16;  // bytes per staged row
  constexpr int kMat = 32 * kStride;
  constexpr int kChunks = HDP / 8;       // 16-byte chunks per staged row
  constexpr int kNC = HDP % 32 == 0 ? 32 : 16;  // output columns per PV pass (divides HDP)
  extern __shared__ __align__(16) uint8_t tsm[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int unit = blockIdx.x * kTWarps + warp;
  if (unit >= p.hw * p.heads) return;
  const int d = unit / p.heads, head = unit - d * p.heads;
  const uint32_t sQ = smem_u32(tsm) + warp * 3 * kMat, sK = sQ + kMat, sV = sK + kMat;

#pragma unroll
  for (int i = 0; i < kChunks; ++i) {
    const int idx = i * 32 + lane;
    const int t = idx / kChunks, c = idx % kChunks;
    const bool col_ok = c * 8 < p.hd;
    const long long row = static_cast<long long>(t) * p.hw + d;
    const uint32_t off = t * kStride + c * 16;
    const bool qok = col_ok && t < p.Tq, kok = col_ok && t < p.Tk;
    cp_async16_zfill(sQ + off, qok ? p.q + row * p.ldq + head * p.hd + c * 8 : p.q, qok ? 16 : 0);
    cp_async16_zfill(sK + off, kok ? p.k + row * p.ldk + head * p.hd + c * 8 : p.k, kok ? 16 : 0);
    cp_async16_zfill(sV + off, kok ? p.v + row * p.ldv + head * p.hd + c * 8 : p.v, kok ? 16 : 0);
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncwarp();

  // ---- S = Q K^T
  float s[2][4][4];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int i = 0; i < 4; ++i) s[mt][nt][i] = 0.f;
#pragma unroll
  for (int ks = 0; ks < HDP / 16; ++ks) {
    uint32_t a[2][4], b[2][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
      ldsm_x4(sQ + (mt * 16 + (lane & 15)) * kStride + (ks * 2 + (lane >> 4)) * 16, a[mt]);
#pragma unroll
    for (int np = 0; np < 2; ++np)  // keys np*16 .. +16: {n-tile 2np: k lo, k hi; n-tile 2np+1: k lo, k hi}
      ldsm_x4(sK + (np * 16 + (lane & 7) + ((lane >> 4) << 3)) * kStride + (ks * 2 + ((lane >> 3) & 1)) * 16, b[np]);
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) mma_bf16_16816(s[mt][nt], a[mt], b[nt >> 1][(nt & 1) * 2], b[nt >> 1][(nt & 1) * 2 + 1]);
  }
  __syncwarp();  // Q rows are dead from here on: they receive the output

  // ---- softmax over keys; this lane holds, per m-tile, rows g = lane/4 (i = 0,1) and g+8 (i = 2,3),
  // key columns nt*8 + (lane%4)*2 + (i&1)
  const float sl2 = p.scale * 1.4426950408889634f;
  uint32_t pa[2][2][4];  // P as A fragments: [m-tile][k-step of 16 keys]
  float inv[2][2];
#pragma unroll
  for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float m = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int col = nt * 8 + (lane & 3) * 2 + i;
          if (col >= p.Tk) s[mt][nt][2 * h + i] = -INFINITY;
          m = fmaxf(m, s[mt][nt][2 * h + i]);
        }
      m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
      m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
      float l = 0.f;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const float e = exp2f((s[mt][nt][2 * h + i] - m) * sl2);
          s[mt][nt][2 * h + i] = e;
          l += e;
        }
      l += __shfl_xor_sync(0xffffffffu, l, 1);
      l += __shfl_xor_sync(0xffffffffu, l, 2);
      inv[mt][h] = 1.f / l;
    }
#pragma unroll
    for (int kk = 0; kk < 2; ++kk) {
      pa[mt][kk][0] = pack_bf16(s[mt][2 * kk][0], s[mt][2 * kk][1]);
      pa[mt][kk][1] = pack_bf16(s[mt][2 * kk][2], s[mt][2 * kk][3]);
      pa[mt][kk][2] = pack_bf16(s[mt][2 * kk + 1][0], s[mt][2 * kk + 1][1]);
      pa[mt][kk][3] = pack_bf16(s[mt][2 * kk + 1][2], s[mt][2 * kk + 1][3]);
    }
  }

  // ---- O = P V, kNC output columns per pass
#pragma unroll 1
  for (int n0 = 0; n0 < HDP; n0 += kNC) {
    float o[2][kNC / 8][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < kNC / 8; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) o[mt][nt][i] = 0.f;
#pragma unroll
    for (int kk = 0; kk < 2; ++kk) {
#pragma unroll
      for (int np = 0; np < kNC / 16; ++np) {
        // matrices: (t lo, n-tile 2np), (t hi, n-tile 2np), (t lo, n-tile 2np+1), (t hi, n-tile 2np+1)
        uint32_t b[4];
        const int mi = lane >> 3;
        ldsm_x4_t(sV + (kk * 16 + (mi & 1) * 8 + (lane & 7)) * kStride + (n0 + np * 16 + (mi >> 1) * 8) * 2, b);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          mma_bf16_16816(o[mt][2 * np], pa[mt][kk], b[0], b[1]);
          mma_bf16_16816(o[mt][2 * np + 1], pa[mt][kk], b[2], b[3]);
        }
      }
    }
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < kNC / 8; ++nt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int row = mt * 16 + (lane >> 2) + 8 * h;
          const uint32_t w = pack_bf16(o[mt][nt][2 * h] * inv[mt][h], o[mt][nt][2 * h + 1] * inv[mt][h]);
          asm volatile("st.shared.b32 [%0], %1;" ::"r"(sQ + row * kStride + (n0 + nt * 8 + (lane & 3) * 2) * 2), "r"(w) : "memory");
        }
  }
  __syncwarp();
#pragma unroll
  for (int i = 0; i < kChunks; ++i) {
    const int idx = i * 32 + lane;
    const int t = idx / kChunks, c = idx % kChunks;
    if (t < p.Tq && c * 8 < p.hd) {
      uint4 w;
      asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w.x), "=r"(w.y), "=r"(w.z), "=r"(w.w) : "r"(sQ + t * kStride + c * 16));
      const long long row = static_cast<long long>(t) * p.hw + d;
      *reinterpret_cast<uint4*>(p.out + row * p.ldo + head * p.hd + c * 8) = w;
    }
  }
}

template <int HDP>
static int launch_temporal_mma(const TemporalParams& p, cudaStream_t stream) {
  constexpr int smem = kTWarps * 3 * 32 * (HDP * 2 + 16);
  static DeviceOnce once;
  int& configured = once.slot();
  if (!configured) {
    int rc = check_cuda(cudaFuncSetAttribute(temporal_attn_mma_kernel<HDP>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                        "temporal mma smem attribute");
    if (rc) return rc;
    configured = 1;
  }
  const int units = p.hw * p.heads;
  return check_cuda(launch_pdl(temporal_attn_mma_kernel<HDP>, dim3((units + kTWarps - 1) / kTWarps), dim3(kTWarps * 32), smem,
                               stream, p),
                    "temporal_attn launch");
}

// ---------------------------------------------------------------------------------------------
// Streaming variant (video_depth_stream.py:134-153, motion_module.py:252-266): one new frame
// attends over Tk <= 32 positions = (Tk-1) cached frames + itself.  The reference re-projects K and
// V of all 32 cached hidden states every frame because the sinusoidal PE is added by *position in
// the current list*, which shifts; projections are linear, so we cache the PE-free projections
// k_c = n W_k^T, v_c = n W_v^T per frame and add the projected PE rows (pe_k[p] = PE[p] W_k^T, ...)
// at read time.  `slots[p]` names the cache slot of list position p; slots[Tk-1] is where the new
// frame's k_c / v_c are stored by this kernel.  Cache layout [hw, cap, C]: the rows of one token's cached frames are
// adjacent (one DRAM page / TLB entry per token instead of one per position).
struct TemporalStreamParams {
  const void* q;     // [hw, ldn] new frame, PE-free projection (bf16, or f32 in the FP32-accuracy mode)
  const void* knew;  // [hw, ldn]
  const void* vnew;
  void* kcache;      // [hw, cap, C], same element type as q
  void* vcache;
  const int* slots;  // device, Tk entries
  const float* pe_q; // [32, C] projected PE tables (fp32)
  const float* pe_k;
  const float* pe_v;
  bf16* out;         // [hw, ldo]; F32: split tensor, parts `split` apart
  int ldn, ldo, hw, heads, hd, C, Tk, split, cap;
  float scale;
};

template <bool F32>
__device__ __forceinline__ void ld8(const void* base, long long off, float (&o)[8]) {
  if (F32) {
    const float4* s = reinterpret_cast<const float4*>(static_cast<const float*>(base) + off);
    const float4 a = s[0], b = s[1];
    o[0] = a.x, o[1] = a.y, o[2] = a.z, o[3] = a.w, o[4] = b.x, o[5] = b.y, o[6] = b.z, o[7] = b.w;
  } else {
    const uint4 v = *reinterpret_cast<const uint4*>(static_cast<const bf16*>(base) + off);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = bf2(w[i]);
      o[2 * i] = f.x, o[2 * i + 1] = f.y;
    }
  }
}
// cache write of 8 values that were read with ld8 (exact: no re-rounding)
template <bool F32>
__device__ __forceinline__ void st8(void* base, long long off, const float (&v)[8]) {
  if (F32) {
    float4* d = reinterpret_cast<float4*>(static_cast<float*>(base) + off);
    d[0] = make_float4(v[0], v[1], v[2], v[3]);
    d[1] = make_float4(v[4], v[5], v[6], v[7]);
  } else {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(static_cast<bf16*>(base) + off) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// The first version of this kernel gave lane p of a warp the K / V rows of cache position p of one (token, head): 32
// lanes reading 16-byte pieces of 32 different cache slots plus three scalar PE loads per element -- LSU-bound, 490 us per
// launch for the 37 x 37, C = 1024 module where the 180 MB of cache it reads are 28 us of HBM time (ncu launch list of
// a streamed vitl frame: 8 launches = 2.3 of the frame's 6 ms).  Here a thread owns one 8-channel chunk of a token row,
// so every cache row is read by consecutive threads as one coalesced run:
//   phase 1  partial dot products of the chunk with the Tk cached K rows (+ projected PE) -> shared memory;
//   phase 2  one warp per (token, head): lane p sums the head's partials of position p, softmax over the positions;
//   phase 3  weighted sum of the Tk cached V rows (+ projected PE) with the same coalesced mapping.
// A thread serves TB tokens with the PE chunk of a position held in registers (the PE tables would otherwise be twice
// the cache traffic).  Block = 128 threads = `groups` token groups of C / 8 threads.
constexpr int kTS2Threads = 128;
template <bool F32>
__device__ __forceinline__ void ldf8(const float* src, float (&o)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(src)), b = __ldg(reinterpret_cast<const float4*>(src + 4));
  o[0] = a.x, o[1] = a.y, o[2] = a.z, o[3] = a.w, o[4] = b.x, o[5] = b.y, o[6] = b.z, o[7] = b.w;
}

// raw 8-element row piece (16 B of bf16 or 32 B of fp32), read-only path; converted after a whole batch has been requested
template <bool F32>
struct Raw8 {
  uint4 a, b;
};
template <bool F32>
__device__ __forceinline__ Raw8<F32> ldraw8(const void* base, long long off) {
  Raw8<F32> r;
  if (F32) {
    const uint4* s = reinterpret_cast<const uint4*>(static_cast<const float*>(base) + off);
    r.a = __ldg(s), r.b = __ldg(s + 1);
  } else {
    r.a = __ldg(reinterpret_cast<const uint4*>(static_cast<const bf16*>(base) + off));
    r.b = make_uint4(0, 0, 0, 0);
  }
  return r;
}
template <bool F32>
__device__ __forceinline__ void cvt8(const Raw8<F32>& r, float (&o)[8]) {
  if (F32) {
    o[0] = __uint_as_float(r.a.x), o[1] = __uint_as_float(r.a.y), o[2] = __uint_as_float(r.a.z), o[3] = __uint_as_float(r.a.w);
    o[4] = __uint_as_float(r.b.x), o[5] = __uint_as_float(r.b.y), o[6] = __uint_as_float(r.b.z), o[7] = __uint_as_float(r.b.w);
  } else {
    const uint32_t w[4] = {r.a.x, r.a.y, r.a.z, r.a.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = bf2(w[i]);
      o[2 * i] = f.x, o[2 * i + 1] = f.y;
    }
  }
}

// The cache rows of JB positions x TB tokens are requested back to back (nothing in the loop stores, so the loads do
// not wait for each other: the first version interleaved the new frame's cache insert with the loop and ran at one
// dependent round trip per position, ~95 us per launch whatever the size); the new frame (position Tk - 1) is handled
// outside the loops from knew / vnew and written to its slot there.
template <bool F32, int TB>
__global__ void __launch_bounds__(kTS2Threads) temporal_attn_stream2_kernel(const TemporalStreamParams p, int groups) {
  pdl_sync();  // launched with programmatic stream serialization (common.cuh)
  extern __shared__ __align__(16) uint8_t tsm[];
#ifndef VDA_TSTREAM_JB
#define VDA_TSTREAM_JB 8
#endif
  constexpr int JB = F32 ? VDA_TSTREAM_JB / 2 : VDA_TSTREAM_JB;
  const int lpt = p.C >> 3;            // threads per token row
  const int used = groups * lpt;
  const int pitch = used + 1;          // row pitch of `part`: phase 2 reads one row per lane (no bank conflicts)
  const int hpl = p.hd >> 3;           // threads (8-channel chunks) per head
  float* part = reinterpret_cast<float*>(tsm);   // [TB][32][pitch]
  float* wgt = part + TB * 32 * pitch;           // [groups * TB * heads][32]
  int* s_slot = reinterpret_cast<int*>(wgt + groups * TB * p.heads * 32);   // [32]
  const int t = threadIdx.x;
  const bool worker = t < used;
  const int g = worker ? t / lpt : 0;
  const int ch = worker ? (t - g * lpt) * 8 : 0;
  const int head = ch / p.hd;
  const int tok0 = (blockIdx.x * groups + g) * TB;
  const int Tk = p.Tk, Tc = p.Tk - 1;   // Tc cached positions + the new frame
  if (t < 32) s_slot[t] = t < Tk ? __ldg(p.slots + t) : 0;
  __syncthreads();
  const int new_slot = s_slot[Tc];

  if (worker) {
    float qv[TB][8];
    {
      float pq[8];
      ldf8<F32>(p.pe_q + static_cast<long long>(Tc) * p.C + ch, pq);
#pragma unroll
      for (int tb = 0; tb < TB; ++tb) {
        const int tok = tok0 + tb;
        if (tok < p.hw) {
          ld8<F32>(p.q, static_cast<long long>(tok) * p.ldn + ch, qv[tb]);
#pragma unroll
          for (int i = 0; i < 8; ++i) qv[tb][i] += pq[i];
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) qv[tb][i] = 0.f;
        }
      }
    }
    for (int j0 = 0; j0 < Tc; j0 += JB) {
      Raw8<F32> raw[JB][TB];
#pragma unroll
      for (int jj = 0; jj < JB; ++jj) {
        const int j = j0 + jj < Tc ? j0 + jj : Tc - 1;   // clamp: the surplus loads repeat the last row
        const int slot = s_slot[j];
#pragma unroll
        for (int tb = 0; tb < TB; ++tb) {
          const int tok = tok0 + tb < p.hw ? tok0 + tb : p.hw - 1;
          raw[jj][tb] = ldraw8<F32>(p.kcache, (static_cast<long long>(tok) * p.cap + slot) * p.C + ch);
        }
      }
      float pe[JB][8];   // the PE chunks of the batch travel with the cache rows (an L2 round trip each otherwise)
#pragma unroll
      for (int jj = 0; jj < JB; ++jj) ldf8<F32>(p.pe_k + static_cast<long long>(j0 + jj < Tc ? j0 + jj : Tc - 1) * p.C + ch, pe[jj]);
#pragma unroll
      for (int jj = 0; jj < JB; ++jj) {
        const int j = j0 + jj;
        if (j < Tc) {
#pragma unroll
          for (int tb = 0; tb < TB; ++tb) {
            float kf[8], dot = 0.f;
            cvt8<F32>(raw[jj][tb], kf);
#pragma unroll
            for (int i = 0; i < 8; ++i) dot = fmaf(qv[tb][i], kf[i] + pe[jj][i], dot);
            part[(tb * 32 + j) * pitch + t] = dot;
          }
        }
      }
    }
    {   // the new frame: position Tc, from knew, inserted into its cache slot
      float pk[8];
      ldf8<F32>(p.pe_k + static_cast<long long>(Tc) * p.C + ch, pk);
#pragma unroll
      for (int tb = 0; tb < TB; ++tb) {
        const int tok = tok0 + tb;
        float dot = 0.f;
        if (tok < p.hw) {
          float kf[8];
          ld8<F32>(p.knew, static_cast<long long>(tok) * p.ldn + ch, kf);
          st8<F32>(p.kcache, (static_cast<long long>(tok) * p.cap + new_slot) * p.C + ch, kf);
#pragma unroll
          for (int i = 0; i < 8; ++i) dot = fmaf(qv[tb][i], kf[i] + pk[i], dot);
        }
        part[(tb * 32 + Tc) * pitch + t] = dot;
      }
    }
  }
  __syncthreads();
  {
    const int warp = t >> 5, lane = t & 31;
    const int pairs = groups * TB * p.heads;
    for (int pr = warp; pr < pairs; pr += kTS2Threads / 32) {
      const int g2 = pr / (TB * p.heads);
      const int rem = pr - g2 * TB * p.heads;
      const int tb = rem / p.heads, h = rem - tb * p.heads;
      float s = -INFINITY;
      if (lane < Tk) {
        const float* src = part + (tb * 32 + lane) * pitch + g2 * lpt + h * hpl;
        float a = 0.f;
        for (int i = 0; i < hpl; ++i) a += src[i];
        s = a * p.scale;
      }
      float m = s;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
      const float e = lane < Tk ? (F32 ? expf(s - m) : __expf(s - m)) : 0.f;
      float l = e;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
      wgt[pr * 32 + lane] = e / l;
    }
  }
  __syncthreads();
  if (!worker) return;
  float acc[TB][8];
#pragma unroll
  for (int tb = 0; tb < TB; ++tb)
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[tb][i] = 0.f;
  const float* wrow = wgt + (g * TB * p.heads + head) * 32;   // + tb * heads * 32 + j
  for (int j0 = 0; j0 < Tc; j0 += JB) {
    Raw8<F32> raw[JB][TB];
#pragma unroll
    for (int jj = 0; jj < JB; ++jj) {
      const int j = j0 + jj < Tc ? j0 + jj : Tc - 1;
      const int slot = s_slot[j];
#pragma unroll
      for (int tb = 0; tb < TB; ++tb) {
        const int tok = tok0 + tb < p.hw ? tok0 + tb : p.hw - 1;
        raw[jj][tb] = ldraw8<F32>(p.vcache, (static_cast<long long>(tok) * p.cap + slot) * p.C + ch);
      }
    }
    float pe[JB][8];
#pragma unroll
    for (int jj = 0; jj < JB; ++jj) ldf8<F32>(p.pe_v + static_cast<long long>(j0 + jj < Tc ? j0 + jj : Tc - 1) * p.C + ch, pe[jj]);
#pragma unroll
    for (int jj = 0; jj < JB; ++jj) {
      const int j = j0 + jj;
      if (j < Tc) {
#pragma unroll
        for (int tb = 0; tb < TB; ++tb) {
          float vf[8];
          cvt8<F32>(raw[jj][tb], vf);
          const float w = wrow[tb * p.heads * 32 + j];
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[tb][i] = fmaf(w, vf[i] + pe[jj][i], acc[tb][i]);
        }
      }
    }
  }
  {
    float pv[8];
    ldf8<F32>(p.pe_v + static_cast<long long>(Tc) * p.C + ch, pv);
#pragma unroll
    for (int tb = 0; tb < TB; ++tb) {
      const int tok = tok0 + tb;
      if (tok < p.hw) {
        float vf[8];
        ld8<F32>(p.vnew, static_cast<long long>(tok) * p.ldn + ch, vf);
        st8<F32>(p.vcache, (static_cast<long long>(tok) * p.cap + new_slot) * p.C + ch, vf);
        const float w = wrow[tb * p.heads * 32 + Tc];
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[tb][i] = fmaf(w, vf[i] + pv[i], acc[tb][i]);
      }
    }
  }
#pragma unroll
  for (int tb = 0; tb < TB; ++tb) {
    const int tok = tok0 + tb;
    if (tok >= p.hw) continue;
    bf16* orow = p.out + static_cast<long long>(tok) * p.ldo + ch;
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h2 = __floats2bfloat162_rn(acc[tb][2 * i], acc[tb][2 * i + 1]);
      hi[i] = *reinterpret_cast<const uint32_t*>(&h2);
      if (F32) {
        const float2 f = __bfloat1622float2(h2);
        const __nv_bfloat162 l2 = __floats2bfloat162_rn(acc[tb][2 * i] - f.x, acc[tb][2 * i + 1] - f.y);
        lo[i] = *reinterpret_cast<const uint32_t*>(&l2);
      }
    }
    *reinterpret_cast<uint4*>(orow) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    if (F32) {
      *reinterpret_cast<uint4*>(orow + p.split) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
      *reinterpret_cast<uint4*>(orow + 2 * p.split) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
    }
  }
}

}  // namespace vda

using namespace vda;

static int launch_temporal_stream(bool f32, const void* q, const void* knew, const void* vnew, int ldn, void* kcache,
                                  void* vcache, const int* slots, const float* pe_q, const float* pe_k, const float* pe_v,
                                  void* out, int ldo, int split, int hw, int heads, int head_dim, int Tk, int cap, void* stream) {
  const int C = heads * head_dim;
  if (!q || !knew || !vnew || !kcache || !vcache || !slots || !pe_q || !pe_k || !pe_v || !out || hw <= 0 || heads <= 0 ||
      head_dim % 8 || C / 8 > kTS2Threads || Tk < 1 || Tk > 32 || cap < Tk || ldn % 8 || ldo % 8 || split < 0 || split % 8)
    return set_error(-1, "vda_temporal_attn_stream: bad arguments (head_dim %% 8, heads * head_dim <= 1024, Tk <= 32 <= cap, ld %% 8)");
  TemporalStreamParams p;
  p.q = q, p.knew = knew, p.vnew = vnew, p.kcache = kcache, p.vcache = vcache;
  p.slots = slots, p.pe_q = pe_q, p.pe_k = pe_k, p.pe_v = pe_v;
  p.out = static_cast<bf16*>(out);
  p.ldn = ldn, p.ldo = ldo, p.hw = hw, p.heads = heads, p.hd = head_dim, p.C = C, p.Tk = Tk, p.split = split, p.cap = cap;
  p.scale = 1.0f / sqrtf(static_cast<float>(head_dim));
  const int lpt = C / 8;
  int groups = kTS2Threads / lpt;
  if (groups > 8) groups = 8;
  const char* etb = getenv("VDA_TSTREAM_TB");   // A/B knob: tokens per thread (2 or 4)
  const int tb = etb && atoi(etb) == 4 ? 4 : 2;   // measured: 2 is faster for every motion-module shape (profiles/r02d_tstream_shapes.log)
  const int used = groups * lpt;
  const int smem = (tb * 32 * (used + 1) + groups * tb * heads * 32 + 32) * 4;
  const int mx = (4 * 32 * (kTS2Threads + 1) + 8 * 4 * 16 * 32 + 32) * 4;   // 129 KB
  if (smem > mx) return set_error(-1, "vda_temporal_attn_stream: too many heads (%d)", heads);
  static DeviceOnce once;
  int& configured = once.slot();
  if (!configured) {
    const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
    int rc = check_cuda(cudaFuncSetAttribute(temporal_attn_stream2_kernel<false, 2>, attr, mx), "temporal stream smem");
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(temporal_attn_stream2_kernel<false, 4>, attr, mx), "temporal stream smem");
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(temporal_attn_stream2_kernel<true, 2>, attr, mx), "temporal stream smem");
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(temporal_attn_stream2_kernel<true, 4>, attr, mx), "temporal stream smem");
    if (rc) return rc;
    configured = 1;
  }
  const dim3 grid((hw + groups * tb - 1) / (groups * tb));
  const cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t le;
  if (f32)
    le = tb == 4 ? launch_pdl(temporal_attn_stream2_kernel<true, 4>, grid, dim3(kTS2Threads), smem, st, p, groups)
                 : launch_pdl(temporal_attn_stream2_kernel<true, 2>, grid, dim3(kTS2Threads), smem, st, p, groups);
  else
    le = tb == 4 ? launch_pdl(temporal_attn_stream2_kernel<false, 4>, grid, dim3(kTS2Threads), smem, st, p, groups)
                 : launch_pdl(temporal_attn_stream2_kernel<false, 2>, grid, dim3(kTS2Threads), smem, st, p, groups);
  return check_cuda(le, "temporal_attn_stream launch");
}

extern "C" int vda_temporal_attn_stream(const void* q, const void* knew, const void* vnew, int ldn, void* kcache,
                                        void* vcache, const int* slots, const float* pe_q, const float* pe_k,
                                        const float* pe_v, void* out, int ldo, int hw, int heads, int head_dim, int Tk,
                                        int cap, void* stream) {
  return launch_temporal_stream(false, q, knew, vnew, ldn, kcache, vcache, slots, pe_q, pe_k, pe_v, out, ldo, 0, hw, heads,
                                head_dim, Tk, cap, stream);
}

extern "C" int vda_temporal_attn_stream_f32(const float* q, const float* knew, const float* vnew, int ldn, float* kcache,
                                            float* vcache, const int* slots, const float* pe_q, const float* pe_k,
                                            const float* pe_v, void* out, int ldo, int split, int hw, int heads,
                                            int head_dim, int Tk, int cap, void* stream) {
  if (split <= 0 || split % 2) return set_error(-1, "vda_temporal_attn_stream_f32: split must be the part stride");
  return launch_temporal_stream(true, q, knew, vnew, ldn, kcache, vcache, slots, pe_q, pe_k, pe_v, out, ldo, split, hw, heads,
                                head_dim, Tk, cap, stream);
}


extern "C" int vda_temporal_attn(const void* q, int ldq, const void* k, int ldk, const void* v, int ldv, void* out,
                                 int ldo, int hw, int heads, int head_dim, int Tq, int Tk, void* stream) {
  if (!q || !k || !v || !out || hw <= 0 || heads <= 0 || head_dim % 8 || head_dim > 128 || Tq < 1 || Tq > 32 || Tk < 1 ||
      Tk > 32 || ldq % 8 || ldk % 8 || ldv % 8 || ldo % 8)
    return set_error(-1, "vda_temporal_attn: bad arguments (head_dim%%8, T<=32, ld%%8)");
  TemporalParams p;
  p.q = static_cast<const bf16*>(q), p.k = static_cast<const bf16*>(k), p.v = static_cast<const bf16*>(v);
  p.out = static_cast<bf16*>(out);
  p.ldq = ldq, p.ldk = ldk, p.ldv = ldv, p.ldo = ldo;
  p.hw = hw, p.heads = heads, p.hd = head_dim, p.Tq = Tq, p.Tk = Tk;
  p.scale = 1.0f / sqrtf(static_cast<float>(head_dim));
  const char* ev = getenv("VDA_TEMPORAL_SIMT");  // A/B knob: the SIMT kernel
  if (!(ev && atoi(ev))) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch ((head_dim + 15) / 16) {
      case 1: return launch_temporal_mma<16>(p, st);
      case 2: return launch_temporal_mma<32>(p, st);
      case 3: return launch_temporal_mma<48>(p, st);
      case 4: return launch_temporal_mma<64>(p, st);
      case 6: return launch_temporal_mma<96>(p, st);
      case 8: return launch_temporal_mma<128>(p, st);
      default: break;  // other head sizes: SIMT kernel below
    }
  }
  const int per_warp = 32 * (head_dim + 2) * 2 + 2 * 32 * head_dim * 2;
  const int smem = per_warp * kTWarps;
  static DeviceOnce once;
  int& configured = once.slot();
  if (smem > configured) {
    int rc = check_cuda(cudaFuncSetAttribute(temporal_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 101376),
                        "temporal smem attribute");
    if (rc) return rc;
    configured = 101376;
  }
  const int units = hw * heads;
  temporal_attn_kernel<<<(units + kTWarps - 1) / kTWarps, kTWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(p);
  return check_cuda(cudaGetLastError(), "temporal_attn launch");
}

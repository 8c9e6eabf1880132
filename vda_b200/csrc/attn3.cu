// Spatial self-attention, third generation: persistent, two independent query-tile streams per CTA whose
// exponential phases alternate ("ping-pong"), one CTA per SM.  Same math and the same TMEM / TMA / tcgen05 data path
// as attn_fwd_v2_kernel (attn.cu); what changes is who runs when.
//
// Why (measured with the phase clocks of the PROF build, scripts/attn_phase_prof.py, vitl block 32 x 16 x S=1370):
// a softmax warp of the v2 kernel spends per KV tile ~1040 clk in its exponentials (MUFU floor 896) and ~900 clk in
// phases that use no MUFU at all (wait for S, tcgen05.ld + row maximum 470, barrier round trips, tcgen05.st 120).
// With two unsynchronised CTAs per SM those phases only overlap by chance: 2500 clk per tile and CTA instead of the
// 2 x 1040 the MUFU pipe allows, and every CTA pays ~5-7 k clk of launch / TMEM allocation / first-load latency /
// epilogue / teardown around its 11 tiles.  Here
//   * a CTA holds TWO streams ("halves"), each a complete copy of the v2 pipeline: producer warp, MMA warp, four
//     softmax warps, its own Q / K / V rings, barriers and 256 TMEM columns (S | P | O);
//   * the softmax warpgroups hand the MUFU pipe to each other with two named barriers: a warpgroup enters its
//     exponentials only when the other one has left its own, so one stream's tcgen05.ld / row maximum / P store /
//     barrier waits always run under the other stream's exponentials;
//   * both streams are persistent over a static round-robin list of (query tile, head, frame) work items, Q double
//     buffered: the next item's Q and first K tiles are in shared memory and its S_0 is computed before the current
//     item's last tile has left the softmax warps, and the epilogue of item i overlaps the first MMAs of item i+1.
// TMEM: 512 columns, allocated once per CTA.
//
// RESULT (B200, profiles/r02d_*attn_v3*): parity-green but SLOWER than v2 in every mode -- 383 us per vitl block with the
// ping-pong, 366 us free-running, 374-384 us chained, against 339 us for v2.  With the exponentials really pinned between
// the barriers (ptxas hoists register-only work above BAR.SYNC; a shared-memory dependency stops it) an exclusive
// exponential phase still takes 1210 clk and each stream's own serial chain (wait S, tcgen05.ld, row maximum, P store,
// item epilogue) is as long as the other stream's exponentials, so strict alternation hides nothing more than chance
// does.  Kept out of the product library: compiled only with -DVDA_ATTN_EXPERIMENTS
// (VDA_NVCC_EXTRA=-DVDA_ATTN_EXPERIMENTS VDA_LIB_NAME=libvda_b200_exp.so python -m vda_b200.build), then selected with
// VDA_ATTN_V=3 (ping-pong; VDA_ATTN_PP=0 free-running) or VDA_ATTN_V=4 (chained, VDA_ATTN_IPC items per CTA).
#ifdef VDA_ATTN_EXPERIMENTS
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "attn_common.cuh"
#include "api_internal.h"

namespace vda {

constexpr int kA3HalfThreads = 192;                 // per stream: warp 0 TMA, warp 1 MMA, warps 2-5 softmax
constexpr int kA3Threads = 2 * kA3HalfThreads;
constexpr int kA3HalfSmem = 2 * kQBytes + 2 * kKBytes + 2 * kVBytes;   // 96 KB
constexpr int kA3Smem = 2 * kA3HalfSmem + 512;
constexpr float kA3RescaleThreshold = 8.0f;

// volatile: the exponentials must stay between the two named barriers of the ping-pong (a plain asm is free to move
// above bar.sync -- its inputs are ready before it -- and ptxas did hoist 31 of the 112 MUFU.EX2 of a tile)
__device__ __forceinline__ float fast_exp2_pinned(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void named_bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int count) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// POLY: one score pair in POLY on the packed polynomial exp2 (0 = none).  PINGPONG = false: the two streams run
// free (A/B comparand).
// GLUED = false ("chained" mode): one stream per CTA, two CTAs per SM as in v2, but a CTA works through p.ipc CONSECUTIVE
// work items with the same cross-item prefetch, so launch / allocation / first-load latency / teardown are paid once
// per p.ipc query tiles while the hardware still hands out CTAs dynamically.
template <int POLY, bool PINGPONG, bool PROF = false, bool GLUED = true>
__global__ void __launch_bounds__(GLUED ? kA3Threads : kA3HalfThreads, GLUED ? 1 : 2)
    attn_fwd_v3_kernel(const __grid_constant__ AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  constexpr int kHalves = GLUED ? 2 : 1;
  const int warp_cta = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int half = GLUED ? warp_cta / 6 : 0, warp = GLUED ? warp_cta % 6 : warp_cta;
  uint8_t* hs = smem + half * kA3HalfSmem;
  uint8_t* sQ = hs;                      // [2][16 KB]
  uint8_t* sK = hs + 2 * kQBytes;        // [2][16 KB]
  uint8_t* sV = sK + 2 * kKBytes;        // [2][16 KB]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kHalves * kA3HalfSmem) + half * 24;
  uint64_t* q_full = bars;         // [2]
  uint64_t* q_empty = bars + 2;    // [2]
  uint64_t* k_full = bars + 4;     // [2]
  uint64_t* k_empty = bars + 6;    // [2]
  uint64_t* v_full = bars + 8;     // [2]
  uint64_t* v_empty = bars + 10;   // [2]
  uint64_t* s_full = bars + 12;
  uint64_t* s_free = bars + 13;
  uint64_t* p_full = bars + 14;
  uint64_t* o_done = bars + 15;
  uint64_t* o_free = bars + 16;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + kHalves * kA3HalfSmem + 2 * 24 * 8);

  // static schedule: slot = 2 * cta + half takes items slot, slot + nslots, ...; item -> (q tile fastest, head, frame)
  // (chained mode: CTA b takes items b * ipc .. b * ipc + ipc - 1)
  const int n_items = p.q_tiles * p.heads * p.frames;
  const int nslots = 2 * gridDim.x;
  const int slot = GLUED ? 2 * blockIdx.x + half : blockIdx.x;
  const int my_items = GLUED ? (slot < n_items ? (n_items - slot + nslots - 1) / nslots : 0)
                             : (n_items - slot * p.ipc < p.ipc ? n_items - slot * p.ipc : p.ipc);
  const int other_slot = 2 * blockIdx.x + (half ^ 1);
  const int other_items = other_slot < n_items ? (n_items - other_slot + nslots - 1) / nslots : 0;
  const int n_tiles = p.kv_tiles;
  const int my_turns = my_items * n_tiles, other_turns = other_items * n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmQK);
    tma_prefetch_desc(&p.tmV);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&q_full[s], 1), mbar_init(&q_empty[s], 1);
      mbar_init(&k_full[s], 1), mbar_init(&k_empty[s], 1);
      mbar_init(&v_full[s], 1), mbar_init(&v_empty[s], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_free, 4);
    mbar_init(p_full, 4);
    mbar_init(o_done, 1);
    mbar_init(o_free, 4);
    fence_mbar_init();
  }
  if (threadIdx.x == 0) tmem_slot[1] = 0u;   // 0.0f, see the ping-pong barrier
  if (warp_cta == 1) {
    tmem_alloc(tmem_slot, 256 * kHalves);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot + half * 256;
  const uint32_t tmem_S = tmem_base, tmem_P = tmem_base + 128, tmem_O = tmem_base + 192;
  pdl_sync();  // the QKV GEMM's outputs are read from here on

  auto item_coords = [&](int it, int& qt, int& head, int& frame) {
    const int i = GLUED ? slot + it * nslots : slot * p.ipc + it;
    qt = i % p.q_tiles;
    const int r = i / p.q_tiles;
    head = r % p.heads;
    frame = r / p.heads;
  };

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    for (int it = 0; it < my_items; ++it) {
      int qt, head, frame;
      item_coords(it, qt, head, frame);
      const int qs = it & 1;
      mbar_wait(&q_empty[qs], ((it >> 1) & 1) ^ 1);
      if (elect_one()) {
        mbar_expect_tx(&q_full[qs], kQBytes);
        tma_load_3d(sQ + qs * kQBytes, &p.tmQK, &q_full[qs], head * 64, qt * 128, frame);
      }
      __syncwarp();
      for (int j = 0; j < n_tiles; ++j) {
        const int g = it * n_tiles + j;
        const int st = g & 1;
        const uint32_t ph = ((g >> 1) & 1) ^ 1;
        mbar_wait(&k_empty[st], ph);
        if (elect_one()) {
          mbar_expect_tx(&k_full[st], kKBytes);
          tma_load_3d(sK + st * kKBytes, &p.tmQK, &k_full[st], p.D + head * 64, j * 128, frame);
        }
        __syncwarp();
        mbar_wait(&v_empty[st], ph);
        if (elect_one()) {
          mbar_expect_tx(&v_full[st], kVBytes);
          tma_load_3d(sV + st * kVBytes, &p.tmV, &v_full[st], j * 128, 0, frame * p.heads + head);
          tma_load_3d(sV + st * kVBytes + kVBytes / 2, &p.tmV, &v_full[st], j * 128 + 64, 0, frame * p.heads + head);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc_s = umma_idesc_f16(128, 128, true);
    constexpr uint32_t idesc_o = umma_idesc_f16(128, 64, true);
    auto issue_s = [&](int g) {   // S_g = Q K^T of global tile g = it * n_tiles + j
      const int it = g / n_tiles, j = g - it * n_tiles;
      const int st = g & 1, qs = it & 1;
      if (j == 0) mbar_wait(&q_full[qs], (it >> 1) & 1);
      mbar_wait(&k_full[st], (g >> 1) & 1);
      if (g > 0) mbar_wait(s_free, (g - 1) & 1);  // softmax has drained S_{g-1} from TMEM
      tc_fence_after();
      const uint64_t qdesc = umma_desc_k_sw128(smem_u32(sQ + qs * kQBytes));
      const uint64_t kdesc = umma_desc_k_sw128(smem_u32(sK + st * kKBytes));
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_S, qdesc + 2 * k, kdesc + 2 * k, idesc_s, k != 0);
        umma_commit(s_full);
        umma_commit(&k_empty[st]);
        if (j == n_tiles - 1) umma_commit(&q_empty[qs]);
      }
      __syncwarp();
    };
    if (my_turns > 0) issue_s(0);
    for (int g = 0; g < my_turns; ++g) {
      const int it = g / n_tiles, j = g - it * n_tiles;
      const int st = g & 1;
      if (g + 1 < my_turns) issue_s(g + 1);
      mbar_wait(&v_full[st], (g >> 1) & 1);
      mbar_wait(p_full, g & 1);
      if (j == 0 && it > 0) mbar_wait(o_free, (it - 1) & 1);   // the previous item's epilogue has read O
      tc_fence_after();
      const uint32_t v_addr = smem_u32(sV + st * kVBytes);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint64_t vdesc = umma_desc_k_sw128(v_addr + (k >> 2) * (64 * 128)) + 2 * (k & 3);
          umma_f16_ts(tmem_O, tmem_P + 8 * k, vdesc, idesc_o, (j | k) != 0);
        }
        umma_commit(o_done);
        umma_commit(&v_empty[st]);
      }
      __syncwarp();
    }
  } else {
    // ------------------------------------------------------------------ softmax warps (thread = query row)
    const int quad = warp_cta & 3;
    const int r = quad * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
    const float sc = p.scale_log2;
    const uint64_t sc2 = pack2(sc, sc);
    // named barriers 1 + h: "stream h may enter its exponentials".  Turn t of stream 0 follows turn t-1 of stream 1,
    // turn t of stream 1 follows turn t of stream 0; a stream only waits / signals while the other one still has turns.
    const int bar_mine = 1 + half, bar_other = 1 + (half ^ 1);
    long long pt[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long long pt_begin = PROF ? clk_after(0.f) : 0;
    for (int it = 0; it < my_items; ++it) {
      int qt, head, frame;
      item_coords(it, qt, head, frame);
      const int q_idx = qt * 128 + r;
      float m_run = -INFINITY, l_run = 0.f;
      for (int j = 0; j < n_tiles; ++j) {
        const int g = it * n_tiles + j;
        long long c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0, c5 = 0;
        if (PROF) c0 = clk_after(l_run);
        mbar_wait(s_full, g & 1);
        tc_fence_after();
        if (PROF) c1 = clk_after(0.f);
        uint32_t s[128];
        {
          uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
          uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
          uint32_t(&s2)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[64]);
          uint32_t(&s3)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[96]);
          tmem_ld32(tmem_S + lane_off + 0, s0);
          tmem_ld32(tmem_S + lane_off + 32, s1);
          tmem_ld32(tmem_S + lane_off + 64, s2);
          tmem_ld32(tmem_S + lane_off + 96, s3);
          tmem_ld_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_free);
        const int kv_left = p.S - j * 128;
        if (kv_left < 128) {
#pragma unroll
          for (int c = 0; c < 128; ++c)
            if (c >= kv_left) s[c] = 0xff800000u;  // -inf
        }
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int c = 0; c < 128; c += 8) {
          mx0 = fmax3(mx0, __uint_as_float(s[c]), __uint_as_float(s[c + 1]));
          mx1 = fmax3(mx1, __uint_as_float(s[c + 2]), __uint_as_float(s[c + 3]));
          mx2 = fmax3(mx2, __uint_as_float(s[c + 4]), __uint_as_float(s[c + 5]));
          mx3 = fmax3(mx3, __uint_as_float(s[c + 6]), __uint_as_float(s[c + 7]));
        }
        const float m_tile = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * sc;
        bool need = false;
        float alpha = 1.f;
        if (j == 0) {
          m_run = m_tile;
        } else {
          need = m_tile > m_run + kA3RescaleThreshold;
          if (need) {
            alpha = fast_exp2(m_run - m_tile);
            m_run = m_tile;
            l_run *= alpha;
          }
        }
        float neg_m = -m_run;
        uint64_t la = pack2(0.f, 0.f), lb = la;
        uint32_t pk[64];
        if (PROF) c2 = clk_after(neg_m);
        if (PINGPONG && GLUED) {
          const bool wait_other = half == 0 ? (g >= 1 && g - 1 < other_turns) : true;   // stream 0 always has >= turns
          if (wait_other) named_bar_sync(bar_mine, 256);
          // ptxas hoists register-only work above BAR.SYNC (31 of a tile's 112 MUFU.EX2 ran before the barrier): make the
          // offset depend on a shared-memory word (always 0.0f) read behind the barrier.
          float z;
          asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(z) : "r"(smem_u32(tmem_slot + 1)) : "memory");
          neg_m += z;
        }
        const uint64_t nm2 = pack2(neg_m, neg_m);
        if (PROF) c3 = clk_after(neg_m);
#pragma unroll
        for (int c0 = 0; c0 < 128; c0 += 32) {
          uint32_t* pc = pk + (c0 >> 1);
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            float x0, x1, p0, p1;
            const uint64_t x2 = ffma2(pack2(__uint_as_float(s[c0 + i]), __uint_as_float(s[c0 + i + 1])), sc2, nm2);
            if (POLY > 0 && ((i >> 1) % POLY) == 0) {   // one pair in POLY entirely on the FMA pipe
              poly_exp2_x2(x2, p0, p1);
            } else {
              unpack2(x2, x0, x1);
              p0 = PINGPONG ? fast_exp2_pinned(x0) : fast_exp2(x0), p1 = PINGPONG ? fast_exp2_pinned(x1) : fast_exp2(x1);
            }
            if (i & 2) lb = fadd2(lb, pack2(p0, p1));
            else la = fadd2(la, pack2(p0, p1));
            __nv_bfloat162 h = __floats2bfloat162_rn(p0, p1);
            pc[i >> 1] = *reinterpret_cast<uint32_t*>(&h);
          }
        }
        if (PROF) {
          float d0, d1;
          unpack2(fadd2(la, lb), d0, d1);
          c4 = clk_after(d0 + d1 + __uint_as_float(pk[63]));
        }
        if (PINGPONG && GLUED) {
          const bool signal_other = half == 0 ? (g < other_turns) : (g + 1 < other_turns);
          if (signal_other) named_bar_arrive(bar_other, 256);
        }
        if (j > 0) {
          // PV_{g-1} reads P and accumulates into O: it must be complete before P is overwritten / O rescaled
          mbar_wait(o_done, (g - 1) & 1);
          tc_fence_after();
          if (__any_sync(0xffffffffu, need)) {
#pragma unroll
            for (int c0 = 0; c0 < 64; c0 += 32) {
              uint32_t o[32];
              tmem_ld32(tmem_O + lane_off + c0, o);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
              tmem_st32(tmem_O + lane_off + c0, o);
            }
          }
        }
#pragma unroll
        for (int c0 = 0; c0 < 128; c0 += 32)
          tmem_st16(tmem_P + lane_off + (c0 >> 1), *reinterpret_cast<uint32_t(*)[16]>(pk + (c0 >> 1)));
        float l0, l1;
        unpack2(fadd2(la, lb), l0, l1);
        l_run += l0 + l1;
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full);
        if (PROF) {
          c5 = clk_after(0.f);
          pt[0] += c1 - c0, pt[1] += c2 - c1, pt[2] += c3 - c2, pt[3] += c4 - c3, pt[4] += c5 - c4;
        }
      }
      const long long ce = PROF ? clk_after(l_run) : 0;
      // ---- epilogue of the item: O / l -> bf16
      mbar_wait(o_done, (it * n_tiles + n_tiles - 1) & 1);
      tc_fence_after();
      uint32_t o0[32], o1[32];
      tmem_ld32(tmem_O + lane_off, o0);
      tmem_ld32(tmem_O + lane_off + 32, o1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(o_free);   // PV_0 of the next item may overwrite O
      const float inv = 1.0f / l_run;
      if (q_idx < p.S) {
        bf16* dst = p.out + (static_cast<long long>(frame) * p.S + q_idx) * p.D + head * 64;
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const uint32_t* o = hh ? o1 : o0;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t w[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(o[8 * c + 2 * i]) * inv,
                                                       __uint_as_float(o[8 * c + 2 * i + 1]) * inv);
              w[i] = *reinterpret_cast<uint32_t*>(&h);
            }
            reinterpret_cast<uint4*>(dst + 32 * hh)[c] = make_uint4(w[0], w[1], w[2], w[3]);
          }
        }
      }
      if (PROF) pt[5] += clk_after(0.f) - ce;
    }
    if (PROF && lane == 0 && my_items > 0) {
      pt[6] = clk_after(0.f) - pt_begin;
      pt[7] = my_items;
      for (int i = 0; i < 8; ++i) p.prof[(static_cast<long long>(slot) * 4 + quad) * 8 + i] = pt[i];
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp_cta == 1) {
    tc_fence_after();
    tmem_dealloc(*tmem_slot, 256 * kHalves);
  }
}

// chained mode: p.ipc > 0
int launch_attn_v3_chained(const AttnParams& p, cudaStream_t st) {
  constexpr int smem = kA3HalfSmem + 512;
  static DeviceOnce once;
  int& configured = once.slot();
  if (!configured) {
    const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
    int rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v3_kernel<8, false, false, false>, attr, smem), "attn3 smem");
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v3_kernel<8, false, true, false>, attr, smem), "attn3 smem");
    if (rc) return rc;
    configured = 1;
  }
  const int n_items = p.q_tiles * p.heads * p.frames;
  const dim3 grid((n_items + p.ipc - 1) / p.ipc);
  const cudaError_t le = p.prof ? launch_pdl(attn_fwd_v3_kernel<8, false, true, false>, grid, dim3(kA3HalfThreads), smem, st, p)
                                : launch_pdl(attn_fwd_v3_kernel<8, false, false, false>, grid, dim3(kA3HalfThreads), smem, st, p);
  return check_cuda(le, "attn3 chained launch");
}

int launch_attn_v3(const AttnParams& p, int poly, bool pingpong, cudaStream_t st) {
  static DeviceOnce once;
  int& configured = once.slot();
  if (!configured) {
    const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
    int rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v3_kernel<8, true>, attr, kA3Smem), "attn3 smem");
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v3_kernel<4, true>, attr, kA3Smem), "attn3 smem");
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v3_kernel<0, true>, attr, kA3Smem), "attn3 smem");
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v3_kernel<8, false>, attr, kA3Smem), "attn3 smem");
    if (rc) return rc;
    configured = 1;
  }
  const int n_items = p.q_tiles * p.heads * p.frames;
  const int sms = num_sms();
  const dim3 grid(static_cast<unsigned>(n_items + 1) / 2 < static_cast<unsigned>(sms) ? (n_items + 1) / 2 : sms);
  cudaError_t le;
  if (p.prof) {
    if (pingpong) {
      check_cuda(cudaFuncSetAttribute(attn_fwd_v3_kernel<8, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kA3Smem), "attn3 smem");
      le = launch_pdl(attn_fwd_v3_kernel<8, true, true>, grid, dim3(kA3Threads), kA3Smem, st, p);
    } else {
      check_cuda(cudaFuncSetAttribute(attn_fwd_v3_kernel<8, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kA3Smem), "attn3 smem");
      le = launch_pdl(attn_fwd_v3_kernel<8, false, true>, grid, dim3(kA3Threads), kA3Smem, st, p);
    }
  } else if (!pingpong) le = launch_pdl(attn_fwd_v3_kernel<8, false>, grid, dim3(kA3Threads), kA3Smem, st, p);
  else if (poly == 8) le = launch_pdl(attn_fwd_v3_kernel<8, true>, grid, dim3(kA3Threads), kA3Smem, st, p);
  else if (poly == 4) le = launch_pdl(attn_fwd_v3_kernel<4, true>, grid, dim3(kA3Threads), kA3Smem, st, p);
  else le = launch_pdl(attn_fwd_v3_kernel<0, true>, grid, dim3(kA3Threads), kA3Smem, st, p);
  return check_cuda(le, "attn3 launch");
}

}  // namespace vda
#endif  // VDA_ATTN_EXPERIMENTS

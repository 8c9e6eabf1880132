// Flash-style spatial self-attention for the DINOv2 encoder on tcgen05 / TMEM / TMA (sm_100a).
//
// Replaces dinov2_layers/attention.py:53-59 (naive softmax(q*scale k^T) v) / :76 (xformers FMHA):
// per (frame, head) non-causal attention over S tokens, head_dim 64, never materialising S x S.
//
// Inputs : qk  [frames*S, 2D] bf16  (Q in cols [0,D), K in cols [D,2D); written by the QKV GEMM)
//          vt  [frames*heads, 64, s_pad] bf16 (V transposed per head, zero padded; same GEMM)
// Output : out [frames*S, D] bf16, head h in cols [64h, 64h+64)
//
// One CTA per (q-tile of 128 rows, head, frame), two CTAs resident per SM so that one CTA's
// exp-bound softmax overlaps the other's MMAs:
//   warp 0    TMA producer: Q once, then K tiles (128x64) and V^T tiles (64x128) through separate 2-stage rings
//   warp 1    TMEM alloc + MMA issuer: S = Q K^T (SS form, M128 N128 K64) and O += P V (TS form: P read from TMEM)
//   warp 2-5  softmax: one thread per query row; the 128-column S row is pulled into registers with one burst of
//             tcgen05.ld, P goes back to TMEM as packed bf16, O stays in TMEM (lazy rescale); see the kernel comment.
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "attn_common.cuh"
#include "api_internal.h"
#include "../../include/vda_b200.h"

namespace vda {

// =============================================================================================
// S row in registers, P and O in TMEM (the first-generation kernel staged P through shared memory and kept O in
// registers: 587 us per vitl block against 340 us; removed).
//   * the softmax thread pulls its whole 128-column S row into registers with one burst of
//     tcgen05.ld and immediately frees the S buffer, so S_{j+1} = Q K_{j+1}^T runs on the tensor pipe
//     while the exponentials of tile j are still being computed;
//   * P_j is written back to TMEM as packed bf16 (tcgen05.st) and the PV MMA takes it as its A operand
//     straight from TMEM (no 32 KB/tile shared-memory round trip, no proxy fence);
//   * O accumulates in TMEM across KV tiles; it is rescaled (tcgen05.ld / st) only when a row
//     maximum grows by more than 2^8 ("lazy rescale": p <= 256, harmless in fp32 / bf16);
//   * K and V have separate 2-stage rings: K_j is released as soon as S_j has been computed.
// TMEM per CTA: S cols [0,128), P cols [128,192), O cols [192,256); two CTAs per SM.
constexpr int kA2Smem = kQBytes + 2 * kKBytes + 2 * kVBytes + 256;
constexpr float kRescaleThreshold = 8.0f;

// POLY: one in POLY exponentials (0 = none, 2/4/8) is evaluated with poly_exp2 on the FMA pipe.
// LATE: the wait for PV_{j-1} (needed before P is overwritten / O rescaled) is taken after the
// exponentials of tile j instead of before them, so that MMA latency hides behind the MUFU work.
// PK: packed f32x2 scale / row-sum and 3-input max (helpers above).
// OPT: optimistic softmax (see the softmax section): no running max / rescale after the first KV tile; a row whose
// sums left the safe range raises flags[cta] and the same kernel compiled with FIXUP recomputes exactly those query
// tiles with the exact online softmax (a second launch that normally finds no flag and exits at once).
template <int POLY, bool LATE, bool PK = false, bool OPT = false, bool FIXUP = false, bool PROF = false>
__global__ void __launch_bounds__(kAttThreads, 2) attn_fwd_v2_kernel(const __grid_constant__ AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int cta_linear = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  if (FIXUP) {
    pdl_wait();
    if (p.flags[cta_linear] == 0) return;   // uniform over the CTA: nothing allocated yet
  }
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  uint8_t* sQ = smem;
  uint8_t* sK = smem + kQBytes;          // [2][16 KB]
  uint8_t* sV = sK + 2 * kKBytes;        // [2][16 KB]  (two 64x64 sub-tiles each)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + 2 * kVBytes);
  uint64_t* q_full = bars;
  uint64_t* k_full = bars + 1;    // [2]
  uint64_t* k_empty = bars + 3;   // [2]
  uint64_t* v_full = bars + 5;    // [2]
  uint64_t* v_empty = bars + 7;   // [2]
  uint64_t* s_full = bars + 9;
  uint64_t* s_free = bars + 10;
  uint64_t* p_full = bars + 11;
  uint64_t* o_done = bars + 12;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = blockIdx.x, head = blockIdx.y, frame = blockIdx.z;
  const int n_tiles = p.kv_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmQK);
    tma_prefetch_desc(&p.tmV);
    mbar_init(q_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&k_full[s], 1), mbar_init(&k_empty[s], 1);
      mbar_init(&v_full[s], 1), mbar_init(&v_empty[s], 1);
    }
    mbar_init(s_full, 1);
    mbar_init(s_free, 4);
    mbar_init(p_full, 4);
    mbar_init(o_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, kAttTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S = tmem_base, tmem_P = tmem_base + 128, tmem_O = tmem_base + 192;
  if (FIXUP) {
    pdl_launch_dependents();
    if (threadIdx.x == 0) p.flags[cta_linear] = 0;   // consumed (every thread has read it before the barrier above)
  } else {
    pdl_sync();  // the QKV GEMM's outputs are read from here on
  }

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      mbar_expect_tx(q_full, kQBytes);
      tma_load_3d(sQ, &p.tmQK, q_full, head * 64, qt * 128, frame);
    }
    __syncwarp();
    for (int j = 0; j < n_tiles; ++j) {
      const int st = j & 1;
      const uint32_t ph = ((j >> 1) & 1) ^ 1;
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
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc_s = umma_idesc_f16(128, 128, true);
    constexpr uint32_t idesc_o = umma_idesc_f16(128, 64, true);
    const uint64_t qdesc = umma_desc_k_sw128(smem_u32(sQ));
    auto issue_s = [&](int j) {
      const int st = j & 1;
      mbar_wait(&k_full[st], (j >> 1) & 1);
      if (j > 0) mbar_wait(s_free, (j - 1) & 1);  // softmax has drained S_{j-1} from TMEM
      tc_fence_after();
      const uint64_t kdesc = umma_desc_k_sw128(smem_u32(sK + st * kKBytes));
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_S, qdesc + 2 * k, kdesc + 2 * k, idesc_s, k != 0);
        umma_commit(s_full);
        umma_commit(&k_empty[st]);
      }
      __syncwarp();
    };
    mbar_wait(q_full, 0);
    issue_s(0);
    for (int j = 0; j < n_tiles; ++j) {
      const int st = j & 1;
      if (j + 1 < n_tiles) issue_s(j + 1);
      mbar_wait(&v_full[st], (j >> 1) & 1);
      mbar_wait(p_full, j & 1);
      tc_fence_after();
      const uint32_t v_addr = smem_u32(sV + st * kVBytes);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          // K = 16 kv positions per MMA: 8 packed-bf16 TMEM columns of P, 32 B of each V^T row
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
    const int quad = warp & 3;
    const int r = quad * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(quad * 32) << 16;
    const int q_idx = qt * 128 + r;
    const float sc = p.scale_log2;
    float m_run = -INFINITY, l_run = 0.f;

    if constexpr (OPT) {
      // ---- optimistic softmax.  The exact online softmax spends a fifth of its instructions and a dependent chain
      // of 16 three-input maxima per KV tile on the running maximum, and its lazy O rescale almost never fires after
      // the first tile.  Here the offset is fixed after tile 0: m_off = (row maximum of tile 0) + 40, so
      // p = 2^(s - m_off) stays finite unless a later score exceeds tile 0's maximum by more than 2^167 (and still
      // resolves weights down to 2^-86 of the row's largest); bf16 P and the fp32 accumulators are scale invariant, so
      // the result is the exact one.  A row whose sum left the safe range (>= 1e30) raises the CTA's flag and the
      // fix-up pass recomputes that query tile with the exact algorithm below.  With no maximum to wait for, the S
      // row is consumed in two halves: the second half's tcgen05.ld is in flight while the first half's exponentials run.
      const uint64_t sc2 = pack2(sc, sc);
      uint64_t la = pack2(0.f, 0.f), lb = la;
      float neg_m = 0.f;
      for (int j = 0; j < n_tiles; ++j) {
        mbar_wait(s_full, j & 1);
        tc_fence_after();
        uint32_t s[128];
        uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
        uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
        uint32_t(&s2)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[64]);
        uint32_t(&s3)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[96]);
        const int kv_left = p.S - j * 128;
        if (j == 0) {
          tmem_ld32(tmem_S + lane_off + 0, s0);
          tmem_ld32(tmem_S + lane_off + 32, s1);
          tmem_ld32(tmem_S + lane_off + 64, s2);
          tmem_ld32(tmem_S + lane_off + 96, s3);
          tmem_ld_wait();
          if (kv_left < 128) {   // a sequence shorter than one tile: its padding must not set the offset
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
          neg_m = -(fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * sc + 40.0f);
        } else {
          tmem_ld32(tmem_S + lane_off + 0, s0);
          tmem_ld32(tmem_S + lane_off + 32, s1);
          tmem_ld_wait();
          tmem_ld32(tmem_S + lane_off + 64, s2);   // in flight during the first half's exponentials
          tmem_ld32(tmem_S + lane_off + 96, s3);
        }
        const uint64_t nm2 = pack2(neg_m, neg_m);
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          if (hh == 1) {
            if (j > 0) tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_free);   // the whole S row is in registers: S_{j+1} may overwrite it
          }
          if (kv_left < 128 && j > 0) {   // last tile: columns beyond the sequence carry no weight (ex2(-inf) = 0)
#pragma unroll
            for (int i = 0; i < 64; ++i)
              if (64 * hh + i >= kv_left) s[64 * hh + i] = 0xff800000u;
          }
          uint32_t pk[32];
#pragma unroll
          for (int i = 0; i < 64; i += 2) {
            const int c = 64 * hh + i;
            float p0, p1;
            const uint64_t x2 = ffma2(pack2(__uint_as_float(s[c]), __uint_as_float(s[c + 1])), sc2, nm2);
            if (POLY > 0 && ((i >> 1) % POLY) == 0) {
              poly_exp2_x2(x2, p0, p1);
            } else {
              float x0, x1;
              unpack2(x2, x0, x1);
              p0 = fast_exp2(x0), p1 = fast_exp2(x1);
            }
            if (i & 2) lb = fadd2(lb, pack2(p0, p1));
            else la = fadd2(la, pack2(p0, p1));
            __nv_bfloat162 h = __floats2bfloat162_rn(p0, p1);
            pk[i >> 1] = *reinterpret_cast<uint32_t*>(&h);
          }
          if (hh == 0 && j > 0) {   // PV_{j-1} reads P: it must be complete before P is overwritten
            mbar_wait(o_done, (j - 1) & 1);
            tc_fence_after();
          }
          tmem_st16(tmem_P + lane_off + 32 * hh, *reinterpret_cast<uint32_t(*)[16]>(&pk[0]));
          tmem_st16(tmem_P + lane_off + 32 * hh + 16, *reinterpret_cast<uint32_t(*)[16]>(&pk[16]));
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full);
      }
      float l0, l1;
      unpack2(fadd2(la, lb), l0, l1);
      l_run = l0 + l1;
      if (__any_sync(0xffffffffu, !(l_run < 1e30f)) && lane == 0) p.flags[cta_linear] = 1;
    } else {
    long long pt[7] = {0, 0, 0, 0, 0, 0, 0};
    const long long pt_begin = PROF ? clk_after(0.f) : 0;
    for (int j = 0; j < n_tiles; ++j) {
      long long c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0, c5 = 0;
      if (PROF) c0 = clk_after(l_run);
      mbar_wait(s_full, j & 1);
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
      if (PROF) c2 = clk_after(__uint_as_float(s[127]));
      const int kv_left = p.S - j * 128;
      if (kv_left < 128) {
#pragma unroll
        for (int c = 0; c < 128; ++c)
          if (c >= kv_left) s[c] = 0xff800000u;  // -inf
      }
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
      if (PK) {
#pragma unroll
        for (int c = 0; c < 128; c += 8) {
          mx0 = fmax3(mx0, __uint_as_float(s[c]), __uint_as_float(s[c + 1]));
          mx1 = fmax3(mx1, __uint_as_float(s[c + 2]), __uint_as_float(s[c + 3]));
          mx2 = fmax3(mx2, __uint_as_float(s[c + 4]), __uint_as_float(s[c + 5]));
          mx3 = fmax3(mx3, __uint_as_float(s[c + 6]), __uint_as_float(s[c + 7]));
        }
      } else {
#pragma unroll
        for (int c = 0; c < 128; c += 4) {
          mx0 = fmaxf(mx0, __uint_as_float(s[c]));
          mx1 = fmaxf(mx1, __uint_as_float(s[c + 1]));
          mx2 = fmaxf(mx2, __uint_as_float(s[c + 2]));
          mx3 = fmaxf(mx3, __uint_as_float(s[c + 3]));
        }
      }
      const float m_tile = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * sc;
      bool need = false;
      float alpha = 1.f;
      if (j == 0) {
        m_run = m_tile;
      } else {
        need = m_tile > m_run + kRescaleThreshold;
        if (need) {
          alpha = fast_exp2(m_run - m_tile);
          m_run = m_tile;
          l_run *= alpha;
        }
        if (!LATE) {
          // PV_{j-1} must be complete before P is overwritten (and before O is rescaled)
          mbar_wait(o_done, (j - 1) & 1);
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
      }
      const float neg_m = -m_run;
      if (PROF) c3 = clk_after(neg_m);
      float l0 = 0.f, l1 = 0.f;
      const uint64_t sc2 = pack2(sc, sc), nm2 = pack2(neg_m, neg_m);
      uint64_t la = pack2(0.f, 0.f), lb = la;
      uint32_t pk[LATE ? 64 : 16];
#pragma unroll
      for (int c0 = 0; c0 < 128; c0 += 32) {
        uint32_t* pc = LATE ? pk + (c0 >> 1) : pk;
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float x0, x1, p0, p1;
          if (PK) {
            const uint64_t x2 = ffma2(pack2(__uint_as_float(s[c0 + i]), __uint_as_float(s[c0 + i + 1])), sc2, nm2);
            if (POLY > 0 && ((i >> 1) % POLY) == 0) {   // one pair in POLY entirely on the FMA pipe
              poly_exp2_x2(x2, p0, p1);
            } else {
              unpack2(x2, x0, x1);
              p0 = fast_exp2(x0), p1 = fast_exp2(x1);
            }
          } else {
            x0 = fmaf(__uint_as_float(s[c0 + i]), sc, neg_m);
            x1 = fmaf(__uint_as_float(s[c0 + i + 1]), sc, neg_m);
            p0 = fast_exp2(x0);
            p1 = (POLY > 0 && ((i >> 1) % (POLY / 2 > 0 ? POLY / 2 : 1)) == 0) ? poly_exp2(x1) : fast_exp2(x1);
          }
          if (PK) {
            if (i & 2) lb = fadd2(lb, pack2(p0, p1));
            else la = fadd2(la, pack2(p0, p1));
          } else {
            l0 += p0;
            l1 += p1;
          }
          __nv_bfloat162 h = __floats2bfloat162_rn(p0, p1);
          pc[i >> 1] = *reinterpret_cast<uint32_t*>(&h);
        }
        if (!LATE) tmem_st16(tmem_P + lane_off + (c0 >> 1), *reinterpret_cast<uint32_t(*)[16]>(pc));
      }
      if (LATE) {
        if (PROF) {
          float d0, d1;
          unpack2(fadd2(la, lb), d0, d1);
          c4 = clk_after(d0 + d1 + __uint_as_float(pk[63]));
        }
        if (j > 0) {
          mbar_wait(o_done, (j - 1) & 1);
          tc_fence_after();
          if (PROF) c5 = clk_after(0.f);
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
      }
      if (PK) unpack2(fadd2(la, lb), l0, l1);
      l_run += l0 + l1;
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
      if (PROF) {
        const long long c6 = clk_after(0.f);
        if (j == 0) c5 = c4;
        pt[0] += c1 - c0, pt[1] += c2 - c1, pt[2] += c3 - c2, pt[3] += c4 - c3, pt[4] += c5 - c4, pt[5] += c6 - c5;
      }
    }
    if (PROF && lane == 0) {
      pt[6] = clk_after(0.f) - pt_begin;
      for (int i = 0; i < 7; ++i) p.prof[(static_cast<long long>(cta_linear) * 4 + quad) * 8 + i] = pt[i];
    }
    }
    // ---- epilogue: O / l -> bf16
    mbar_wait(o_done, (n_tiles - 1) & 1);
    tc_fence_after();
    const float inv = 1.0f / l_run;
    bf16* dst = p.out + (static_cast<long long>(frame) * p.S + q_idx) * p.D + head * 64;
#pragma unroll
    for (int c0 = 0; c0 < 64; c0 += 32) {
      uint32_t o[32];
      tmem_ld32(tmem_O + lane_off + c0, o);
      tmem_ld_wait();
      if (q_idx < p.S) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t w[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(o[8 * c + 2 * i]) * inv,
                                                     __uint_as_float(o[8 * c + 2 * i + 1]) * inv);
            w[i] = *reinterpret_cast<uint32_t*>(&h);
          }
          reinterpret_cast<uint4*>(dst + c0)[c] = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kAttTmemCols);
  }
}

}  // namespace vda

#ifdef VDA_ATTN_EXPERIMENTS   // A/B builds only (see attn3.cu): persistent / ping-pong variants and the phase clocks
namespace vda {
int launch_attn_v3(const AttnParams& p, int poly, bool pingpong, cudaStream_t st);   // attn3.cu
int launch_attn_v3_chained(const AttnParams& p, cudaStream_t st);
}
#endif
using namespace vda;

#ifdef VDA_ATTN_EXPERIMENTS
// Phase-clock instrumentation (scripts/attn_phase_prof.py): not declared in the public header.
static long long* g_attn_prof = nullptr;
extern "C" void vda_debug_attn_prof(long long* buf) { g_attn_prof = buf; }
#endif

extern "C" int vda_flash_attn_fwd(const void* qk, const void* vt, void* out, int frames, int S, int heads, int s_pad,
                                  int* fixup_flags, void* stream) {
  if (!qk || !vt || !out || frames <= 0 || S <= 0 || heads <= 0 || s_pad < S || s_pad % 8)
    return set_error(-1, "vda_flash_attn_fwd: bad arguments");
  const int D = heads * 64;
  AttnParams p;
  memset(&p, 0, sizeof(p));
  {
    const uint64_t dims[3] = {(uint64_t)(2 * D), (uint64_t)S, (uint64_t)frames};
    const uint64_t str[2] = {(uint64_t)(2 * D) * 2, (uint64_t)S * 2 * D * 2};
    const uint32_t box[3] = {64, 128, 1};
    int rc = encode_bf16_map(&p.tmQK, qk, 3, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)s_pad, 64, (uint64_t)frames * heads};
    const uint64_t str[2] = {(uint64_t)s_pad * 2, (uint64_t)s_pad * 64 * 2};
    const uint32_t box[3] = {64, 64, 1};
    int rc = encode_bf16_map(&p.tmV, vt, 3, dims, str, box);
    if (rc) return rc;
  }
  p.out = static_cast<bf16*>(out);
  p.S = S, p.D = D, p.heads = heads, p.frames = frames;
  p.q_tiles = (S + 127) / 128;
  p.kv_tiles = (S + 127) / 128;
  p.scale_log2 = 0.125f * 1.4426950408889634f;
  p.flags = fixup_flags;
#ifdef VDA_ATTN_EXPERIMENTS
  p.prof = g_attn_prof;
#endif
  // A/B knobs: VDA_ATTN_POLY = 0 / 4 / 8 (default): one score pair in N evaluated with the packed polynomial exp2;
  // VDA_ATTN_OPT = 1 (experiment builds, -DVDA_ATTN_EXPERIMENTS, only; the product library always runs the exact online
  // softmax and leaves `fixup_flags` untouched): optimistic softmax + fix-up pass when the caller passes flags.  Measured on B200
  // it is 5 % SLOWER in the window (10.5 -> 11.0 ms, profiles/r02c_ab_attn_optimistic_negative.log) although it issues a
  // third fewer instructions per KV tile and has no maximum chain.  What the phase clocks and pipe probes of round 2 show
  // (profiles/r02d_attn_v2_phase_clocks_*.log, r02d_pipe_rate_all.log, r02d_ncu_attn_phase_buckets.txt): TMEM reads are
  // NOT a limit (tcgen05.ld sustains ~900 B/clk/SM; a 64 KB S tile is ~150 clk of latency); a softmax warp alone on its
  // scheduler needs ~1040 clk per KV tile for the exponentials (MUFU floor 896 with 1 pair in 8 on the FMA pipe) plus
  // ~900 clk in which it issues no MUFU at all (wait for S incl. the CTA's first TMA round trip, tcgen05.ld + row maximum
  // ~470, barrier round trips, P store, epilogue).  Two co-resident CTAs overlap those phases by chance: 2500 clk per tile
  // and CTA = 1250 per SM against the 900-1040 the MUFU pipe would allow.
  const char* ep = getenv("VDA_ATTN_POLY");
  const int poly = ep ? atoi(ep) : 8;
#ifdef VDA_ATTN_EXPERIMENTS
  const char* eo = getenv("VDA_ATTN_OPT");
  const bool opt = fixup_flags != nullptr && eo && atoi(eo) == 1;
  // VDA_ATTN_V = 3: persistent ping-pong kernel (attn3.cu; VDA_ATTN_PP = 0: free-running streams); 4: chained mode
  // (VDA_ATTN_IPC work items per CTA); default 2 = this file.  All measured slower (profiles/r02d_ab_attn_v3_*).
  const char* ev = getenv("VDA_ATTN_V");
  const int version = ev ? atoi(ev) : 2;
  if (version == 4 && !opt) {
    const char* ei = getenv("VDA_ATTN_IPC");
    p.ipc = ei ? atoi(ei) : 2;
    if (p.ipc < 1) p.ipc = 1;
    return launch_attn_v3_chained(p, static_cast<cudaStream_t>(stream));
  }
  if (version == 3 && !opt) {
    const char* epp = getenv("VDA_ATTN_PP");
    return launch_attn_v3(p, poly, !(epp && atoi(epp) == 0), static_cast<cudaStream_t>(stream));
  }
#endif
  static DeviceOnce once;
  int& configured = once.slot();
  if (!configured) {
    const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
    int rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v2_kernel<8, true, true>, attr, kA2Smem), "attn smem");
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v2_kernel<4, true, true>, attr, kA2Smem), "attn smem");
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v2_kernel<0, true, true>, attr, kA2Smem), "attn smem");
#ifdef VDA_ATTN_EXPERIMENTS
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v2_kernel<8, true, true, true>, attr, kA2Smem), "attn smem");
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v2_kernel<0, true, true, true>, attr, kA2Smem), "attn smem");
    if (!rc) rc = check_cuda(cudaFuncSetAttribute(attn_fwd_v2_kernel<8, true, true, false, true>, attr, kA2Smem), "attn smem");
#endif
    if (rc) return rc;
    configured = 1;
  }
  const dim3 grid(p.q_tiles, heads, frames);
  const cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t le;
#ifdef VDA_ATTN_EXPERIMENTS
  if (p.prof) {
    const char* e1 = getenv("VDA_ATTN_ONE_CTA");   // phase clocks without a co-resident CTA: pad the smem request
    const int smem = (e1 && atoi(e1) == 1) ? 120 * 1024 : kA2Smem;
    check_cuda(cudaFuncSetAttribute(attn_fwd_v2_kernel<8, true, true, false, false, true>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, smem), "attn smem");
    le = launch_pdl(attn_fwd_v2_kernel<8, true, true, false, false, true>, grid, dim3(kAttThreads), smem, st, p);
  } else if (opt) {
    le = poly == 0 ? launch_pdl(attn_fwd_v2_kernel<0, true, true, true>, grid, dim3(kAttThreads), kA2Smem, st, p)
                   : launch_pdl(attn_fwd_v2_kernel<8, true, true, true>, grid, dim3(kAttThreads), kA2Smem, st, p);
    if (le == cudaSuccess)   // fix-up pass: exact kernel on the flagged query tiles only (normally none)
      le = launch_pdl(attn_fwd_v2_kernel<8, true, true, false, true>, grid, dim3(kAttThreads), kA2Smem, st, p);
  } else
#endif
  if (poly == 8) {
    le = launch_pdl(attn_fwd_v2_kernel<8, true, true>, grid, dim3(kAttThreads), kA2Smem, st, p);
  } else if (poly == 4) {
    le = launch_pdl(attn_fwd_v2_kernel<4, true, true>, grid, dim3(kAttThreads), kA2Smem, st, p);
  } else {
    le = launch_pdl(attn_fwd_v2_kernel<0, true, true>, grid, dim3(kAttThreads), kA2Smem, st, p);
  }
  return check_cuda(le, "attn launch");
}

// tcgen05 / TMEM / TMA GEMM and 3x3 implicit-GEMM kernels for sm_100a.
//
//   C[M,N] = epilogue( A[M,K] * W[N,K]^T )      A, W bf16 (K-major), fp32 accumulate in TMEM
//
// Replaces every nn.Linear / 1x1 conv / 3x3 conv / ConvTranspose(k=s) call site of the reference
// hot path: dinov2_layers/attention.py:51,60, mlp.py:36-39, dpt.py:60-124,
// util/blocks.py:78-91,160, motion_module/motion_module.py:115,128, motion_module/attention.py:81-83.
//
// Two persistent kernels share one fused epilogue (gemm_epilogue.cuh); 12 warps per CTA:
//   warp 0    TMA producer    (128B-swizzled K-major tiles, mbarrier ring)
//   warp 1    MMA issuer      (one thread)
//   warp 2    TMEM allocator  (2 accumulator stages of BN fp32 columns)
//   warp 4+   epilogue        (16 warps: 4 lane quadrants x 4 column slices; tcgen05.ld 32x32b ->
//                              registers -> fused epilogue -> global)
//
//   gemm_tc2_kernel<BN>  CTA pairs (cluster of 2, tcgen05.mma.cta_group::2, M=256 x N=BN per pair):
//                        each CTA loads its 128 rows of A and HALF of the W tile, so the operand
//                        bytes pulled from L2 per FLOP drop by a third against the 1-CTA tile and a
//                        stage is 32 KB instead of 48 KB (6-8 stages in flight).  Used whenever the
//                        problem has enough 256-row tile pairs to fill the 74 SM pairs.
//   gemm_tc_kernel<BN>   single CTA, M=128 x N=BN: small problems (streaming mode, vits head).
#include "gemm_epilogue.cuh"
#include "api_internal.h"

namespace vda {

constexpr int kGemmThreads = 128 + kEpiThreads;  // 4 control warps + epilogue warps
constexpr int kRowOffBytes = 2 * kBlockM * 8;     // epilogue MODE 3: output row offsets x 2 tile parities

template <int BN>
struct GemmCfg {
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = BN * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN >= 256) ? 4 : (BN >= 128 ? 6 : 8);
  static constexpr int kTmemCols = (2 * BN <= 32) ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));
  static constexpr int kBiasBytes = 2 * 2 * BN * 4;  // {bias, scale} x 2 tile parities
  // halo mode (3x3 convs, BN <= 64) re-partitions the same buffer into stages of {patch, 9 W tiles}
  static constexpr int kHaloStageBytes = 23 * 1024 + 9 * BN * kBlockK * 2;
  static constexpr int kHaloStages = BN <= 32 ? 3 : (BN <= 64 ? 2 : 0);
  static constexpr int kRingBytes = (kStages * kStageBytes > kHaloStages * kHaloStageBytes) ? kStages * kStageBytes : kHaloStages * kHaloStageBytes;
  static constexpr int kSmemBytes = kRingBytes + 1024 /*align slack*/ + 256 /*barriers*/ + kBiasBytes + kRowOffBytes;
};

template <int BN>
struct Gemm2Cfg {
  static constexpr int kABytes = kBlockM * kBlockK * 2;
  static constexpr int kBBytes = (BN / 2) * kBlockK * 2;  // this CTA's half of the W tile
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN >= 256) ? 6 : 8;
  static constexpr int kTmemCols = GemmCfg<BN>::kTmemCols;
  static constexpr int kBiasBytes = 2 * 2 * BN * 4;
  static constexpr int kBarBytes = 512;   // up to kMaxRing2 full + empty barriers, 2 + 2 accumulator, 2 + 2 halo patch
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + kBarBytes + kBiasBytes + kRowOffBytes;
};
constexpr int kMaxRing2 = 16;

// halo mode of the 1-CTA kernel: patch of an 8 x 16-pixel tile plus a 1-pixel border, one 64-channel block
constexpr int kHaloW = 10, kHaloH = 18;
constexpr int kHaloBytes = kHaloW * kHaloH * 128;                  // 23,040 B written by one TMA box
constexpr int kHaloSlotBytes = (kHaloBytes + 1023) / 1024 * 1024;  // keeps the W tiles behind it 1024-aligned

struct TileCoord {
  int img, h0, w0;
};
__device__ __forceinline__ TileCoord conv_tile_coord(const GemmParams& p, int m_t) {
  TileCoord c;
  const int tpi = p.tiles_w * p.tiles_h;
  c.img = m_t / tpi;
  const int t = m_t - c.img * tpi;
  const int ty = t / p.tiles_w, tx = t - ty * p.tiles_w;
  c.h0 = ty * (kBlockM >> p.tw_log2);
  c.w0 = tx << p.tw_log2;
  return c;
}

// ------------------------------------------------------------------------------------ 1-CTA kernel
template <int BN>
__global__ void __launch_bounds__(kGemmThreads, 1) gemm_tc_kernel(const __grid_constant__ GemmParams p) {
  using Cfg = GemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kRingBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  float* s_epi = reinterpret_cast<float*>(smem + Cfg::kRingBytes + 256);
  // bytes of the operand ring this launch uses; the epilogue's boxes / residual ring live behind it
  const int ring_used = p.stages * (p.halo ? Cfg::kHaloStageBytes : Cfg::kStageBytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.m_tiles * p.n_tiles;
  // Tile order (tile = m_tile * n_tiles + n_tile).  If the A row-blocks of all CTAs fit in L2 together
  // (K <= ~1k), each CTA takes a CONTIGUOUS run of tiles: its A rows come from HBM once, every later
  // n_tile hits L2, and the next row-block is prefetched into L2 a tile ahead, so first-touch HBM
  // latency never reaches the smem ring.  Otherwise tiles are dealt round-robin: the tiles in flight
  // then form a narrow band of the (m, n) grid (few A row-blocks x all W tiles) that stays L2-resident.
  const bool chunked = p.sched == 2;  // measured: round-robin wins on every encoder shape, contiguous runs kept as a knob
  const int chunk = (num_tiles + gridDim.x - 1) / gridDim.x;
  const int tile_begin = chunked ? blockIdx.x * chunk : blockIdx.x;
  const int tile_end = chunked ? min(num_tiles, tile_begin + chunk) : num_tiles;
  const int tile_step = chunked ? 1 : gridDim.x;

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
  pdl_sync();  // everything above overlapped the previous kernel's tail; global memory from here on

  // ---- halo mode (3x3 convs with few output channels).  Fetching every (tap, channel block) A tile
  // separately pulls each input pixel through L2 nine times; with BN <= 64 there is too little math
  // per byte to hide that (Cout = 32 tail: 2.5 ms at ~8 TB/s of L2 traffic).  Here a stage holds the
  // 10 x 18-pixel halo patch of one channel block (ONE TMA box, 23 KB instead of 9 x 16 KB) plus the W
  // tiles of its 9 taps, and the taps are nine tap-shifted UMMA descriptors into the patch: row 8g+i of
  // the A operand is pixel (y=g, x=i) of the 8 x 16 tile, i.e. patch row (g+dy)*10 + i+dx -> start
  // address + (dy*10+dx)*128, group pitch 1280 B.  36 MMAs are issued per barrier wait.
  if (p.halo && warp == 0) {
    const int cblocks = p.C / kBlockK;
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = tile_begin; tile < tile_end; tile += tile_step) {
      const int m_t = tile / p.n_tiles, n_t = tile - m_t * p.n_tiles;
      const TileCoord tc = conv_tile_coord(p, m_t);
      for (int cb = 0; cb < cblocks; ++cb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* st = smem + stage * Cfg::kHaloStageBytes;
        if (elect_one()) {
          mbar_expect_tx(&full_bar[stage], kHaloBytes + 9 * Cfg::kBBytes);
          tma_load_4d(st, &p.tmH, &full_bar[stage], cb * kBlockK, tc.w0 - 1, tc.h0 - 1, tc.img);
#pragma unroll
          for (int tap = 0; tap < 9; ++tap)
            tma_load_2d(st + kHaloSlotBytes + tap * Cfg::kBBytes, &p.tmB, &full_bar[stage], (tap * cblocks + cb) * kBlockK, n_t * BN);
        }
        __syncwarp();
        if (++stage == p.stages) stage = 0, phase ^= 1;
      }
    }
  } else if (p.halo && warp == 1) {
    constexpr uint32_t idesc = umma_idesc_f16(kBlockM, BN, true);
    const int cblocks = p.C / kBlockK;
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = tile_begin; tile < tile_end; tile += tile_step, ++it) {
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(&tempty_bar[as], aphase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      for (int cb = 0; cb < cblocks; ++cb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t patch = smem_u32(smem + stage * Cfg::kHaloStageBytes);
        if (elect_one()) {
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            const int dy = tap / 3, dx = tap - dy * 3;
            const uint64_t adesc = umma_desc_k_sw128_sbo(patch + (dy * kHaloW + dx) * 128, kHaloW * 128);
            const uint64_t bdesc = umma_desc_k_sw128(patch + kHaloSlotBytes + tap * Cfg::kBBytes);
#pragma unroll
            for (int k = 0; k < kBlockK / 16; ++k) umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (cb | tap | k) != 0);
          }
          umma_commit(&empty_bar[stage]);
          if (cb == cblocks - 1) umma_commit(&tfull_bar[as]);
        }
        __syncwarp();
        if (++stage == p.stages) stage = 0, phase ^= 1;
      }
    }
  } else
  // The producer and MMA loops are executed by their WHOLE warp (all addresses and descriptors stay
  // warp-uniform, i.e. in uniform registers) and only the issuing instructions are predicated on one
  // elected lane.  Running the loops under `if (lane == 0)` made ptxas wrap every tcgen05.mma in an
  // elect / R2UR.BROADCAST sequence and the single issuing thread, not the tensor pipe, set the pace.
  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    const int cblocks = (p.a_mode == A_CONV3X3) ? (p.C / kBlockK) : 1;
    for (int tile = tile_begin; tile < tile_end; tile += tile_step) {
      const int m_t = tile / p.n_tiles, n_t = tile - m_t * p.n_tiles;
      TileCoord tc = {0, 0, 0};
      if (p.a_mode == A_CONV3X3) tc = conv_tile_coord(p, m_t);
      if (chunked && n_t == p.n_tiles - 1 && tile + 1 < tile_end && elect_one()) {
        for (int kb = 0; kb < p.num_k_blocks; ++kb) tma_prefetch_2d(&p.tmA, kb * kBlockK, (m_t + 1) * kBlockM);
      }
      __syncwarp();
      for (int kb = 0; kb < p.num_k_blocks; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * Cfg::kStageBytes;
        uint8_t* sb = sa + Cfg::kABytes;
        if (elect_one()) {
          mbar_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          if (p.a_mode == A_CONV3X3) {
            const int tap = kb / cblocks, cb = kb - tap * cblocks;
            const int kh = tap / 3, kw = tap - kh * 3;
            tma_load_4d(sa, &p.tmA, &full_bar[stage], cb * kBlockK, tc.w0 + kw - 1, tc.h0 + kh - 1, tc.img);
          } else {
            tma_load_2d(sa, &p.tmA, &full_bar[stage], kb * kBlockK, m_t * kBlockM);
          }
          tma_load_2d(sb, &p.tmB, &full_bar[stage], kb * kBlockK, n_t * BN);
        }
        __syncwarp();
        if (++stage == p.stages) stage = 0, phase ^= 1;
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_f16(kBlockM, BN, true);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = tile_begin; tile < tile_end; tile += tile_step, ++it) {
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
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            // +32 B per K=16 step inside the 128 B swizzle row (start-address field is >>4)
            umma_f16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          }
          umma_commit(&empty_bar[stage]);
          if (kb == p.num_k_blocks - 1) umma_commit(&tfull_bar[as]);
        }
        __syncwarp();
        if (++stage == p.stages) stage = 0, phase ^= 1;
      }
    }
  } else if (warp >= 4) {
    const int ew = warp - 4;
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    const int half = ew >> 2;  // column slice of this warp (0..kEpiParts-1)
    const int r = quad * 32 + lane;
    const int et = threadIdx.x - 128;
    const TileSched ts = {tile_begin, tile_step, tile_end, 1, 0};
    ResRing rr = make_res_ring<BN>(p, ts, smem + ring_used, Cfg::kRingBytes - ring_used, et, r, half);
    // box epilogues: epi_boxes(mode) boxes per column slice in the freed stages; thread 0 of each slice issues the TMA
    const uint32_t red_base = smem_u32(smem + ring_used) + static_cast<uint32_t>(half) * epi_boxes(p.epi_tma) * kRedBoxBytes;
    const bool red_issuer = (et & 127) == 0;
    int box_ctr = 0;
    int it = 0;
    for (int tile = tile_begin; tile < tile_end; tile += tile_step, ++it) {
      const int m_t = tile / p.n_tiles, n_t = tile - m_t * p.n_tiles;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      float* s_bias = s_epi + as * 2 * BN;
      float* s_scale = s_bias + BN;
      long long* s_rowoff = reinterpret_cast<long long*>(s_epi + 4 * BN) + as * kBlockM;
      stage_row_offsets(p, m_t, et, s_rowoff);
      stage_bias_scale<BN>(p, n_t, et, s_bias, s_scale);
      if (p.epi_tma == 5) coop_tile_begin<BN>(p, n_t, r, half, red_base, box_ctr, s_rowoff);
      if (tile + tile_step < tile_end)  // residual rows of the NEXT tile into L2: a whole tile of lead time
        epilogue_prefetch<BN>(p, (tile + tile_step) / p.n_tiles, (tile + tile_step) % p.n_tiles, r, half);
      mbar_wait(&tfull_bar[as], aphase);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + as * BN;
      // QKV tiles that hold V columns keep the transposing element-wise store
      const bool v_tile = p.store_mode == ST_QKV && (n_t + 1) * BN > 2 * p.qkv_d;
      if (p.epi_tma == 1)
        epilogue_tile_box<BN, 1>(p, m_t, n_t, t_acc, r, half, s_bias, s_scale, red_base, box_ctr, red_issuer);
      else if (p.epi_tma == 2 && !v_tile)
        epilogue_tile_box<BN, 2>(p, m_t, n_t, t_acc, r, half, s_bias, s_scale, red_base, box_ctr, red_issuer);
      else if (p.epi_tma == 3)
        epilogue_tile_box<BN, 3>(p, m_t, n_t, t_acc, r, half, s_bias, s_scale, red_base, box_ctr, red_issuer, s_rowoff);
      else if (p.epi_tma == 4)
        epilogue_tile_box<BN, 4>(p, m_t, n_t, t_acc, r, half, s_bias, s_scale, red_base, box_ctr, red_issuer);
      else if (p.epi_tma == 5)
        epilogue_tile_coop<BN>(p, n_t, t_acc, r, half, s_bias, s_scale, red_base, box_ctr, s_rowoff);
      else
        epilogue_tile<BN>(p, m_t, n_t, t_acc, r, half, s_bias, s_scale, ts, rr, it);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);
    }
    if (p.epi_tma && p.epi_tma != 5 && red_issuer) bulk_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------ 2-CTA kernel
// Cluster of two CTAs = one SM pair.  The pair owns a 256 x BN output tile: CTA r holds rows
// [128r, 128r+128) (its own TMEM lanes) and feeds W rows [r*BN/2, (r+1)*BN/2) of the tile.  Only
// the leader (rank 0) issues MMAs; its tcgen05.commit is multicast so the "slot free" and
// "accumulator full" barriers of BOTH CTAs flip; both CTAs' TMA loads credit the leader's full barrier;
// both CTAs' epilogue warps arrive on the leader's "accumulator empty" barrier.
template <int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
    gemm_tc2_kernel(const __grid_constant__ GemmParams p) {
  using Cfg = Gemm2Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  // the dynamic smem base offset is identical in both CTAs, so aligned offsets match across the pair
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* empty_bar = full_bar + kMaxRing2;
  uint64_t* tfull_bar = empty_bar + kMaxRing2;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* pfull_bar = tempty_bar + 2;   // halo mode: the two patch slots
  uint64_t* pempty_bar = pfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(pempty_bar + 2);
  float* s_epi = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes + Cfg::kBarBytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int m2_tiles = (p.m_tiles + 1) >> 1;
  const int num_tiles = m2_tiles * p.n_tiles;
  // tile order: see gemm_tc_kernel
  const bool chunked = p.sched == 2;  // measured: round-robin wins on every encoder shape, contiguous runs kept as a knob
  const int chunk = (num_tiles + num_clusters - 1) / num_clusters;
  const int tile_begin = chunked ? cluster_id * chunk : cluster_id;
  const int tile_end = chunked ? min(num_tiles, tile_begin + chunk) : num_tiles;
  const int tile_step = chunked ? 1 : num_clusters;
  // work units: tiles, with the tiles of the partial last wave cut into ksplit K ranges (round-robin only)
  const int unit_end = (p.ksplit > 1 && !chunked) ? num_tiles + (num_tiles - p.tail_begin) * (p.ksplit - 1) : tile_end;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kMaxRing2; ++s) {
      mbar_init(&full_bar[s], 1);   // leader's: one arrive.expect_tx by the leader's producer
      mbar_init(&empty_bar[s], 1);  // one multicast commit from the leader's MMA thread
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);               // multicast commit
      mbar_init(&tempty_bar[s], 2 * kEpiWarps);  // leader's: epilogue warps of both CTAs
      mbar_init(&pfull_bar[s], 1);
      mbar_init(&pempty_bar[s], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_2cta(tmem_slot, Cfg::kTmemCols);
    tmem_relinquish_2cta();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_sync();  // everything above overlapped the previous kernel's tail; global memory from here on

  // ---- halo mode of the CTA-pair kernel (3x3 convs; see the 1-CTA kernel for the idea).  With nine separate
  // (tap, channel block) A tiles a 256 x 128 pair tile pulls 96 B/clk/SM of operands through L2 -> SM, far above
  // what the fabric sustains (output_conv1 at 296x296: tensor pipe 46 % active, 19.9 GB of L2 -> SM traffic for a
  // 1.4 GB map).  Here each CTA fetches ONE 10 x 18-pixel halo patch per 64-channel block (two slots) and the nine
  // taps are tap-shifted UMMA descriptors into it; the W tiles (this CTA's half, one per tap) stream through their
  // own ring of p.stages slots behind the patches.  Operand traffic per channel block: 23 + 9 x 8 KB instead of
  // 9 x 24 KB (BN = 128).
  constexpr int kWBytes = Cfg::kBBytes;
  uint8_t* const w_ring = smem + 2 * kHaloSlotBytes;
  if (p.halo && warp == 0) {
    const int cblocks = p.C / kBlockK;
    int ws = 0, ps = 0;
    uint32_t wphase = 0, pphase = 0;
    for (int v = tile_begin; v < tile_end; v += tile_step) {
      const int m2 = v / p.n_tiles, n_t = v - m2 * p.n_tiles;
      const TileCoord tc = conv_tile_coord(p, 2 * m2 + static_cast<int>(rank));
      for (int cb = 0; cb < cblocks; ++cb) {
        mbar_wait(&pempty_bar[ps], pphase ^ 1);
        if (elect_one()) {
          if (leader) mbar_expect_tx(&pfull_bar[ps], 2 * kHaloBytes);
          tma_load_4d_2sm(smem + ps * kHaloSlotBytes, &p.tmH, &pfull_bar[ps], cb * kBlockK, tc.w0 - 1, tc.h0 - 1, tc.img);
        }
        __syncwarp();
        if (++ps == 2) ps = 0, pphase ^= 1;
        for (int tap = 0; tap < 9; ++tap) {
          mbar_wait(&empty_bar[ws], wphase ^ 1);
          if (elect_one()) {
            if (leader) mbar_expect_tx(&full_bar[ws], 2 * kWBytes);
            tma_load_2d_2sm(w_ring + ws * kWBytes, &p.tmB, &full_bar[ws], (tap * cblocks + cb) * kBlockK,
                            n_t * BN + static_cast<int>(rank) * (BN / 2));
          }
          __syncwarp();
          if (++ws == p.stages) ws = 0, wphase ^= 1;
        }
      }
    }
  } else if (p.halo && warp == 1) {
    if (leader) {
      constexpr uint32_t idesc = umma_idesc_f16(2 * kBlockM, BN, true);
      const int cblocks = p.C / kBlockK;
      int ws = 0, ps = 0, it = 0;
      uint32_t wphase = 0, pphase = 0;
      for (int v = tile_begin; v < tile_end; v += tile_step, ++it) {
        const int as = it & 1;
        mbar_wait(&tempty_bar[as], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int cb = 0; cb < cblocks; ++cb) {
          mbar_wait(&pfull_bar[ps], pphase);
          const uint32_t patch = smem_u32(smem + ps * kHaloSlotBytes);
          for (int tap = 0; tap < 9; ++tap) {
            mbar_wait(&full_bar[ws], wphase);
            tc_fence_after();
            const int dy = tap / 3, dx = tap - dy * 3;
            const uint64_t adesc = umma_desc_k_sw128_sbo(patch + (dy * kHaloW + dx) * 128, kHaloW * 128);
            const uint64_t bdesc = umma_desc_k_sw128(smem_u32(w_ring + ws * kWBytes));
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < kBlockK / 16; ++k)
                umma_f16_ss_2cta(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (cb | tap | k) != 0);
              umma_commit_2cta(&empty_bar[ws]);
              if (tap == 8) {
                umma_commit_2cta(&pempty_bar[ps]);
                if (cb == cblocks - 1) umma_commit_2cta(&tfull_bar[as]);
              }
            }
            __syncwarp();
            if (++ws == p.stages) ws = 0, wphase ^= 1;
          }
          if (++ps == 2) ps = 0, pphase ^= 1;
        }
      }
    }
  } else
  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    const int cblocks = (p.a_mode == A_CONV3X3) ? (p.C / kBlockK) : 1;
    for (int v = tile_begin; v < unit_end; v += tile_step) {
      const WorkUnit u = decode_unit(p, v);
      const int tile = u.tile;
      const int m2 = tile / p.n_tiles, n_t = tile - m2 * p.n_tiles;
      const int m_t = 2 * m2 + static_cast<int>(rank);
      TileCoord tc = {0, 0, 0};
      if (p.a_mode == A_CONV3X3) tc = conv_tile_coord(p, m_t);
      if (chunked && n_t == p.n_tiles - 1 && tile + 1 < tile_end && elect_one()) {
        for (int kb = 0; kb < p.num_k_blocks; ++kb) tma_prefetch_2d(&p.tmA, kb * kBlockK, (m_t + 2) * kBlockM);
      }
      __syncwarp();
      for (int kb = u.kb0; kb < u.kb1; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * Cfg::kStageBytes;
        uint8_t* sb = sa + Cfg::kABytes;
        if (elect_one()) {
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * Cfg::kStageBytes);
          if (p.a_mode == A_CONV3X3) {
            const int tap = kb / cblocks, cb = kb - tap * cblocks;
            const int kh = tap / 3, kw = tap - kh * 3;
            tma_load_4d_2sm(sa, &p.tmA, &full_bar[stage], cb * kBlockK, tc.w0 + kw - 1, tc.h0 + kh - 1, tc.img);
          } else {
            tma_load_2d_2sm(sa, &p.tmA, &full_bar[stage], kb * kBlockK, m_t * kBlockM);
          }
          tma_load_2d_2sm(sb, &p.tmB, &full_bar[stage], kb * kBlockK, n_t * BN + static_cast<int>(rank) * (BN / 2));
        }
        __syncwarp();
        if (++stage == p.stages) stage = 0, phase ^= 1;
      }
    }
  } else if (warp == 1) {
    if (leader) {
      constexpr uint32_t idesc = umma_idesc_f16(2 * kBlockM, BN, true);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int v = tile_begin; v < unit_end; v += tile_step, ++it) {
        const WorkUnit u = decode_unit(p, v);
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tempty_bar[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        for (int kb = u.kb0; kb < u.kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint64_t adesc = umma_desc_k_sw128(sa);
          const uint64_t bdesc = umma_desc_k_sw128(sa + Cfg::kABytes);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < kBlockK / 16; ++k)
              umma_f16_ss_2cta(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, ((kb - u.kb0) | k) != 0);
            umma_commit_2cta(&empty_bar[stage]);
            if (kb == u.kb1 - 1) umma_commit_2cta(&tfull_bar[as]);
          }
          __syncwarp();
          if (++stage == p.stages) stage = 0, phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    const int ew = warp - 4;
    const int quad = warp & 3;
    const int half = ew >> 2;  // column slice of this warp (0..kEpiParts-1)
    const int r = quad * 32 + lane;
    const int et = threadIdx.x - 128;
    const TileSched ts = {tile_begin, tile_step, tile_end, 2, static_cast<int>(rank)};
    // operand bytes this launch uses; the epilogue's boxes / residual ring live behind them
    const int ring_used = p.halo ? 2 * kHaloSlotBytes + p.stages * kWBytes : p.stages * Cfg::kStageBytes;
    ResRing rr = make_res_ring<BN>(p, ts, smem + ring_used, Cfg::kStages * Cfg::kStageBytes - ring_used, et, r, half);
    // box epilogues: epi_boxes(mode) boxes per column slice in the freed stages; thread 0 of each slice issues the TMA
    const uint32_t red_base = smem_u32(smem + ring_used) + static_cast<uint32_t>(half) * epi_boxes(p.epi_tma) * kRedBoxBytes;
    const bool red_issuer = (et & 127) == 0;
    int box_ctr = 0;
    int it = 0;
    for (int v = tile_begin; v < unit_end; v += tile_step, ++it) {
      const WorkUnit u = decode_unit(p, v);
      const int tile = u.tile;
      const int m2 = tile / p.n_tiles, n_t = tile - m2 * p.n_tiles;
      const int m_t = 2 * m2 + static_cast<int>(rank);
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      float* s_bias = s_epi + as * 2 * BN;
      float* s_scale = s_bias + BN;
      long long* s_rowoff = reinterpret_cast<long long*>(s_epi + 4 * BN) + as * kBlockM;
      stage_row_offsets(p, m_t, et, s_rowoff);
      stage_bias_scale<BN>(p, n_t, et, s_bias, s_scale, u.kb0 == 0);  // the bias rides with the first K range only
      if (p.epi_tma == 5) coop_tile_begin<BN>(p, n_t, r, half, red_base, box_ctr, s_rowoff);
      if (tile + tile_step < tile_end)
        epilogue_prefetch<BN>(p, 2 * ((tile + tile_step) / p.n_tiles) + static_cast<int>(rank),
                              (tile + tile_step) % p.n_tiles, r, half);
      mbar_wait(&tfull_bar[as], aphase);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + as * BN;
      // QKV tiles that hold V columns keep the transposing element-wise store
      const bool v_tile = p.store_mode == ST_QKV && (n_t + 1) * BN > 2 * p.qkv_d;
      if (p.epi_tma == 1)
        epilogue_tile_box<BN, 1>(p, m_t, n_t, t_acc, r, half, s_bias, s_scale, red_base, box_ctr, red_issuer);
      else if (p.epi_tma == 2 && !v_tile)
        epilogue_tile_box<BN, 2>(p, m_t, n_t, t_acc, r, half, s_bias, s_scale, red_base, box_ctr, red_issuer);
      else if (p.epi_tma == 3)
        epilogue_tile_box<BN, 3>(p, m_t, n_t, t_acc, r, half, s_bias, s_scale, red_base, box_ctr, red_issuer, s_rowoff);
      else if (p.epi_tma == 4)
        epilogue_tile_box<BN, 4>(p, m_t, n_t, t_acc, r, half, s_bias, s_scale, red_base, box_ctr, red_issuer);
      else if (p.epi_tma == 5)
        epilogue_tile_coop<BN>(p, n_t, t_acc, r, half, s_bias, s_scale, red_base, box_ctr, s_rowoff);
      else
        epilogue_tile<BN>(p, m_t, n_t, t_acc, r, half, s_bias, s_scale, ts, rr, it);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tempty_bar[as], 0);
    }
    if (p.epi_tma && p.epi_tma != 5 && red_issuer) bulk_wait_all();
  }

  // nobody may exit (or free TMEM) while the peer can still touch this CTA's smem, barriers or TMEM
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2cta(tmem_base, Cfg::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------ launch
template <int BN>
static cudaError_t launch_bn(const GemmParams& p, int num_sms, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  static DeviceOnce once;
  int& configured = once.slot();
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    if (e != cudaSuccess) return e;
    configured = 1;
  }
  const int tiles = p.m_tiles * p.n_tiles;
  const int grid = tiles < num_sms ? tiles : num_sms;
  return launch_pdl(gemm_tc_kernel<BN>, dim3(grid), dim3(kGemmThreads), Cfg::kSmemBytes, stream, p);
}

template <int BN>
static cudaError_t launch2_bn(const GemmParams& p, int num_sms, cudaStream_t stream) {
  using Cfg = Gemm2Cfg<BN>;
  static DeviceOnce once;
  int& configured = once.slot();
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc2_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
    if (e != cudaSuccess) return e;
    configured = 1;
  }
  int pairs = ((p.m_tiles + 1) / 2) * p.n_tiles;
  if (p.ksplit > 1) pairs += (pairs - p.tail_begin) * (p.ksplit - 1);   // work units: K ranges of the split tiles
  const int max_pairs = num_sms / 2;
  const int grid = 2 * (pairs < max_pairs ? pairs : max_pairs);
  return launch_pdl(gemm_tc2_kernel<BN>, dim3(grid), dim3(kGemmThreads), Cfg::kSmemBytes, stream, p);  // cluster dims are compiled in
}

cudaError_t launch_gemm(const GemmParams& p, int block_n, bool two_cta, int num_sms, cudaStream_t stream) {
  if (two_cta) {
    switch (block_n) {
      case 64: return launch2_bn<64>(p, num_sms, stream);
      case 128: return launch2_bn<128>(p, num_sms, stream);
      case 256: return launch2_bn<256>(p, num_sms, stream);
      default: return cudaErrorInvalidValue;
    }
  }
  switch (block_n) {
    case 32: return launch_bn<32>(p, num_sms, stream);
    case 64: return launch_bn<64>(p, num_sms, stream);
    case 128: return launch_bn<128>(p, num_sms, stream);
    case 256: return launch_bn<256>(p, num_sms, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace vda

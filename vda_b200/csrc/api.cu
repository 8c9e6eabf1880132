// C ABI (include/vda_b200.h): argument checking, TMA descriptor encoding, kernel launches.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/vda_b200.h"
#include "gemm_params.h"
#include "api_internal.h"

namespace vda {
cudaError_t launch_gemm(const GemmParams& p, int block_n, bool two_cta, int num_sms, cudaStream_t stream);

static thread_local char g_err[512] = "";
static int g_sms[kMaxDevices] = {};  // per device, filled by vda_init(device)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  return set_error(static_cast<int>(e), "%s: %s", what, cudaGetErrorString(e));
}
int current_device() {
  int d = 0;
  cudaGetDevice(&d);
  return d;
}
int num_sms() { return g_sms[current_device() & (kMaxDevices - 1)]; }
// Programmatic dependent launch (common.cuh).  Measured on B200: inside the CUDA graph of a 32-frame window
// it is 2 % SLOWER (596 -> 584 frames/s, two interleaved runs each), for the ~300 single-tile launches of a
// streaming frame 2 % faster (2.215 -> 2.168 ms), so the host turns it on per launch sequence
// (vda_set_pdl, default off); VDA_PDL=0/1 overrides both ways for A/B runs.
static int g_pdl = 0;
bool pdl_enabled() {
  static const int forced = [] {
    const char* v = getenv("VDA_PDL");
    return v ? (atoi(v) != 0 ? 1 : 0) : -1;
  }();
  return forced >= 0 ? forced != 0 : g_pdl != 0;
}

static int encode_map(CUtensorMap* tm, CUtensorMapDataType dt, CUtensorMapSwizzle sw, const void* base, int rank,
                      const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box);

// bf16 tensor map, 128B swizzle, inner box = 64 elements (128 B); rank 2..4.
int encode_bf16_map(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                    const uint32_t* box) {
  return encode_map(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, CU_TENSOR_MAP_SWIZZLE_128B, base, rank, dims, strides_bytes, box);
}

// bf16 tensor map without swizzle (boxes that threads, not the UMMA, read back)
int encode_bf16_map_plain(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                          const uint32_t* box) {
  return encode_map(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, CU_TENSOR_MAP_SWIZZLE_NONE, base, rank, dims, strides_bytes, box);
}

static int encode_map(CUtensorMap* tm, CUtensorMapDataType dt, CUtensorMapSwizzle sw, const void* base, int rank,
                      const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box) {
  if (!g_encode) return set_error(-2, "vda_init() has not been called");
  cuuint64_t gdim[5], gstr[5];
  cuuint32_t bdim[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    estr[i] = 1;
    if (i > 0) {
      gstr[i - 1] = strides_bytes[i - 1];
      if (gstr[i - 1] % 16) return set_error(-3, "tensor map stride %llu not a multiple of 16 B", (unsigned long long)gstr[i - 1]);
    }
  }
  if (reinterpret_cast<uintptr_t>(base) % 16) return set_error(-3, "tensor map base not 16 B aligned");
  CUresult r = g_encode(tm, dt, rank, const_cast<void*>(base), gdim, gstr, bdim, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(-4, "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
  return 0;
}

// Tile selection.  CTA pairs (256 x BN per pair, half of W per CTA) whenever some BN in {256,128,64}
// wastes <= 12.5 % of the padded N and leaves at least one full wave of pair tiles for the 74 SM
// pairs; otherwise the single-CTA kernel with the largest BN that still fills the SMs.
static void pick_tile(const vda_epilogue* e, int N, int m_tiles, int* bn_out, bool* two_out) {
  if (e->store_mode == ST_HEADDOT) {
    *bn_out = 32, *two_out = false;
    return;
  }
  const int n16 = ((N + 15) / 16) * 16;
  if (e->cta_pair != 1) {
    const int m2 = (m_tiles + 1) / 2;
    const int cands2[3] = {256, 128, 64};
    for (int i = 0; i < 3; ++i) {
      const int bn = cands2[i];
      if (e->block_n && e->block_n != bn) continue;
      const int nt = (N + bn - 1) / bn;
      const bool forced = e->cta_pair == 2;
      if (!forced && (nt * bn - n16) * 8 > nt * bn) continue;
      // at least (nearly) one full wave of pair tiles: 72 tiles on 74 pairs (streaming qkv, M = 1370, N = 3072) is one
      // wave of 256-wide tiles instead of two waves of 128-wide ones.  VDA_GEMM_WAVE_FRAC8 = 8 restores "one full wave".
      static const int frac8 = [] { const char* e = getenv("VDA_GEMM_WAVE_FRAC8"); return e ? atoi(e) : 7; }();
      if (!forced && static_cast<long long>(m2) * nt * 8 < static_cast<long long>(num_sms() / 2) * frac8) continue;
      *bn_out = bn, *two_out = true;
      return;
    }
  }
  *two_out = false;
  if (e->block_n) {
    *bn_out = e->block_n;
    return;
  }
  // single-CTA kernel: the BN with the cheapest schedule, cost = waves x (BN + a fixed per-tile share).  "Largest BN
  // that still fills the SMs" left e.g. the streaming fc1 (M = 1370, N = 4096) with 176 tiles of 256 columns on 148
  // SMs = two full-length waves, the second one 19 % occupied; 352 tiles of 128 columns take three half-length waves.
  const int cands[4] = {256, 128, 64, 32};
  long long best_cost = -1;
  *bn_out = N >= 64 ? 64 : 32;
  for (int i = 0; i < 4; ++i) {
    const int bn = cands[i];
    if (bn > 32 && bn > ((N + 31) / 32) * 32) continue;
    const long long tiles = static_cast<long long>(m_tiles) * ((N + bn - 1) / bn);
    const long long waves = (tiles + num_sms() - 1) / num_sms();
    const long long cost = waves * (bn + 48);
    if (best_cost < 0 || cost < best_cost) best_cost = cost, *bn_out = bn;
  }
}

// Small-M problems whose epilogue is the in-place fp32 residual update (proj / fc2 of a streaming frame: M = 1370):
// too few tiles for the SMs but a long K.  Every 256 x 256 pair tile is cut into `ks` K ranges handled by different CTA
// pairs whose partial sums meet in the residual through the TMA reduce-add (the tail split-K mechanism with
// tail_begin = 0).  Returns ks >= 2, or 0 when the launch does not qualify.
static int small_m_ksplit(const vda_epilogue* e, int M, int N, int K) {
  const char* kv = getenv("VDA_GEMM_KSPLIT");  // A/B knob: 0 disables
  if (kv && atoi(kv) == 0) return 0;
  if (e->cta_pair == 1 || (e->block_n && e->block_n != 256) || N % 256 || e->act != ACT_NONE) return 0;
  if (!e->res_f32 || e->res_f32 != e->out_f32 || e->out || e->out_relu || e->res1 || e->res2 || e->split ||
      e->store_mode != ST_PLAIN)
    return 0;
  const int pairs = num_sms() / 2, tiles = ((M + 255) / 256) * (N / 256), kblocks = (K + 63) / 64;
  if (tiles >= pairs) return 0;
  int ks = pairs / tiles;
  if (ks > kblocks / 4) ks = kblocks / 4;
  if (ks > 8) ks = 8;
  return ks >= 2 ? ks : 0;
}

// Epilogue flavour (gemm_epilogue.cuh): 1 = in-place fp32 residual update as TMA reduce-add,
// 2 = bf16 output through TMA box stores, 0 = generic element-wise path.
static int pick_epi_tma(const vda_epilogue* e, bool conv) {
  const char* ev = getenv("VDA_GEMM_NO_TMA_EPI");  // A/B knob: 1 disables all, 2 disables the bf16 store flavour
  const int off_lvl = ev ? atoi(ev) : 0;
  const bool off = off_lvl == 1;
  const auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  if (!off && !e->split && (e->out_relu || e->res1 || e->res2)) {
    // bf16 residuals / ReLU'd second output: cooperative 64-byte-row boxes (gemm_epilogue.cuh, MODE 5)
    const bool ok = e->store_mode == ST_PLAIN && e->act != ACT_GEGLU && !e->out_f32 && !e->res_f32 && (e->out || e->out_relu) &&
                    e->ldo % 8 == 0 && al16(e->out) && al16(e->out_relu) && al16(e->res1) && al16(e->res2) && off_lvl != 5;
    return ok ? 5 : 0;
  }
  if (off || e->out_relu || e->res1 || e->res2 || e->split) return 0;
  if (e->act == ACT_GEGLU) {  // value * gelu(gate) pairs -> bf16 boxes of 32 outputs (64 accumulator columns)
    const bool ok = !conv && e->store_mode == ST_PLAIN && e->out && !e->out_f32 && !e->res_f32 && e->ldo % 8 == 0 &&
                    (reinterpret_cast<uintptr_t>(e->out) & 15) == 0 && off_lvl != 4;
    return ok ? 4 : 0;
  }
  const auto aligned = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
  const int ldf = e->ld32 ? e->ld32 : e->ldo;
  if (!conv && e->store_mode == ST_PLAIN && e->out_f32 && !e->out && ldf % 4 == 0 && aligned(e->out_f32) &&
      (e->res_f32 == e->out_f32 || !e->res_f32))
    return 1;  // in-place residual update (reduce-add) or plain fp32 output (store)
  if ((e->store_mode == ST_PLAIN || (!conv && e->store_mode == ST_QKV)) && e->out && !e->out_f32 && !e->res_f32 &&
      e->ldo % 8 == 0 && aligned(e->out) && off_lvl != 2)
    return 2;
  // ConvTranspose pixel-shuffle store: box staging + cooperative scattered store (4 lanes per 64-byte row run)
  if (!conv && e->store_mode == ST_CONVT && e->out && !e->out_f32 && !e->res_f32 && e->ct_cout % 32 == 0 && e->ldo % 8 == 0 &&
      aligned(e->out) && off_lvl != 3)
    return 3;
  return 0;
}

// smem ring depth of a launch: the compiled maximum of the kernel variant, minus the stages whose
// shared memory the epilogue takes over (48 KB: TMA boxes or the fp32-residual staging ring).
static int pick_stages(int bn, bool two, const vda_epilogue* e, int epi_tma) {
  const int max_stages = two ? (bn >= 256 ? 6 : 8) : (bn >= 256 ? 4 : (bn >= 128 ? 6 : 8));
  const bool staged = epi_tma || (e->res_f32 && e->store_mode == ST_PLAIN && e->act != ACT_GEGLU);
  if (!staged) return max_stages;
  const int stage_bytes = 128 * 64 * 2 + (two ? bn / 2 : bn) * 64 * 2;
  const int need = epi_tma ? epi_box_area_bytes(epi_tma) : 48 * 1024;  // TMA boxes, or the 3-4 slot fp32-residual staging ring
  const int freed = (need + stage_bytes - 1) / stage_bytes;
  return max_stages - freed;
}

static int fill_epilogue(GemmParams& p, const vda_epilogue* e) {
  if (num_sms() <= 0) return set_error(-2, "vda_init(device) has not been called for the current device %d", current_device());
  const char* sv = getenv("VDA_GEMM_SCHED");  // tuning knob: 1 = round-robin tiles, 2 = contiguous runs
  p.sched = sv ? atoi(sv) : 0;
  p.act = e->act;
  p.store_mode = e->store_mode;
  p.bias = e->bias;
  p.scale = e->scale;
  p.res_f32 = e->res_f32;
  p.res1 = static_cast<const __nv_bfloat16*>(e->res1);
  p.res2 = static_cast<const __nv_bfloat16*>(e->res2);
  p.out = static_cast<__nv_bfloat16*>(e->out);
  p.out_relu = static_cast<__nv_bfloat16*>(e->out_relu);
  p.out_f32 = e->out_f32;
  p.ldo = e->ldo;
  p.n_out = (e->act == ACT_GEGLU) ? p.N / 2 : p.N;
  p.ct_s = e->ct_s, p.ct_cout = e->ct_cout, p.ct_h = e->ct_h, p.ct_w = e->ct_w;
  p.vt = static_cast<__nv_bfloat16*>(e->vt);
  p.qkv_d = e->qkv_d, p.qkv_heads = e->qkv_heads, p.qkv_s = e->qkv_s, p.qkv_spad = e->qkv_spad;
  p.w2 = e->w2;
  p.b2 = e->b2;
  p.split = e->split;
  p.ld32 = e->ld32;
  if (e->ld32 % 4) return set_error(-1, "ld32 must be a multiple of 4");
  if (e->split && (e->split % 8 || e->store_mode == ST_QKV)) return set_error(-1, "split must be a multiple of 8; not with the QKV store");
  if (p.N % 16) return set_error(-1, "N=%d must be a multiple of 16", p.N);
  if (e->act == ACT_GEGLU) {
    if (p.N % 32 || !e->bias || !e->out) return set_error(-1, "GEGLU needs N%%32==0, bias and out");
    if (e->store_mode != ST_PLAIN) return set_error(-1, "GEGLU supports plain stores only");
  }
  if (e->store_mode == ST_HEADDOT) {
    if (p.N > 32 || !e->bias || !e->w2 || !e->out_f32) return set_error(-1, "HEADDOT needs N<=32, bias, w2, out_f32");
  } else {
    if (e->ldo % 8) return set_error(-1, "ldo=%d must be a multiple of 8", e->ldo);
    if (!e->out && !e->out_f32 && !e->out_relu && e->store_mode != ST_QKV) return set_error(-1, "no output given");
  }
  if (e->store_mode == ST_CONVT && (e->ct_cout % 16 || e->ct_s * e->ct_s * e->ct_cout != p.N))
    return set_error(-1, "CONVT needs ct_cout%%16==0 and N == s*s*cout");
  if (e->store_mode == ST_QKV && (!e->vt || 3 * e->qkv_d != p.N || e->qkv_heads * 64 != e->qkv_d))
    return set_error(-1, "QKV store needs vt, N == 3*d and head_dim 64");
  return 0;
}

}  // namespace vda

using namespace vda;

extern "C" {

int vda_init(int device) {
  int rc = check_cuda(cudaSetDevice(device), "cudaSetDevice");
  if (rc) return rc;
  rc = check_cuda(cudaFree(0), "cudaFree(0)");
  if (rc) return rc;
  cudaDeviceProp prop;
  rc = check_cuda(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties");
  if (rc) return rc;
  if (prop.major != 10) return set_error(-5, "device %d is sm_%d%d; this library is sm_100a only", device, prop.major, prop.minor);
  if (device < 0 || device >= kMaxDevices) return set_error(-5, "device index %d out of range", device);
  g_sms[device] = prop.multiProcessorCount;
  if (!g_encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    rc = check_cuda(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q), "cudaGetDriverEntryPoint");
    if (rc) return rc;
    if (!fn || q != cudaDriverEntryPointSuccess) return set_error(-6, "cuTensorMapEncodeTiled not available");
    g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  }
  return 0;
}

const char* vda_last_error(void) { return g_err; }
int vda_num_sms(void) { return num_sms(); }
#ifndef VDA_SRC_DIGEST
#define VDA_SRC_DIGEST "unstamped"
#endif
/* sha256 of csrc/ + include/ + the nvcc flags this binary was built from (vda_b200/build.py); the marker prefix lets
 * the build script read it out of the file without loading it */
const char* vda_build_digest(void) { return &"vda-src-digest:" VDA_SRC_DIGEST[15]; }
int vda_set_pdl(int on) {
  const int prev = g_pdl;
  g_pdl = on != 0;
  return prev;
}

int vda_gemm_bf16(const void* A, int lda, const void* W, int ldw, int M, int N, int K, const vda_epilogue* epi,
                  void* stream) {
  if (!A || !W || !epi || M <= 0 || N <= 0 || K <= 0) return set_error(-1, "vda_gemm_bf16: bad arguments");
  if (lda % 8 || ldw % 8) return set_error(-1, "vda_gemm_bf16: lda/ldw must be multiples of 8");
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = M, p.N = N, p.K = K;
  p.a_mode = A_PLAIN;
  p.num_k_blocks = (K + 63) / 64;
  p.m_tiles = (M + 127) / 128;
  int rc = fill_epilogue(p, epi);
  if (rc) return rc;
  int bn;
  bool two;
  pick_tile(epi, N, p.m_tiles, &bn, &two);
  p.epi_tma = pick_epi_tma(epi, false);
  const int small_ks = p.epi_tma == 1 ? small_m_ksplit(epi, M, N, K) : 0;
  if (small_ks) bn = 256, two = true;
  p.n_tiles = (N + bn - 1) / bn;
  if (p.epi_tma == 4 && bn < 128) p.epi_tma = 0;  // a GEGLU box spans 64 accumulator columns of one column slice
  if (p.epi_tma == 5 && N % 32) p.epi_tma = 0;    // whole 32-column boxes only
  p.stages = pick_stages(bn, two, epi, p.epi_tma);
  if (p.epi_tma == 1) {
    const char* rs = getenv("VDA_GEMM_RED_AS_STORE");  // knob for timing experiments only (drops the residual)
    p.f32_store = epi->res_f32 == nullptr || (rs && atoi(rs));
    const uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
    const uint64_t str[1] = {(uint64_t)(epi->ld32 ? epi->ld32 : epi->ldo) * 4};
    const uint32_t box[2] = {16, 128};
    if ((rc = encode_map(&p.tmD, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, CU_TENSOR_MAP_SWIZZLE_64B, epi->out_f32, 2, dims, str, box)))
      return rc;
    // Tail split-K: proj / fc2 of the encoder are 688 pair tiles for 74 SM pairs = 9.3 waves, i.e. the
    // tenth wave keeps 22 pairs busy and 52 idle (7 % of the launch).  Their epilogue is a reduce-add into
    // the fp32 residual, so the R tiles of the partial wave are cut into ksplit K ranges (one per otherwise
    // idle pair) whose partial products simply meet in the residual; the bias goes with the first range.
    // fp32 adds commute, so only the rounding order of those few tiles varies from run to run.
    const char* kv = getenv("VDA_GEMM_KSPLIT");  // A/B knob: 0 disables
    if (two && !p.f32_store && epi->act == ACT_NONE && p.sched != 2 && !(kv && atoi(kv) == 0)) {
      const int tiles = ((p.m_tiles + 1) / 2) * p.n_tiles, nc = num_sms() / 2;
      const int rem = tiles % nc;
      if (tiles > nc && rem) {
        int ks = nc / rem;
        if (ks > p.num_k_blocks / 4) ks = p.num_k_blocks / 4;
        if (ks > 8) ks = 8;
        if (ks >= 2) p.ksplit = ks, p.tail_begin = tiles - rem;
      }
    }
    if (small_ks && !p.f32_store && p.sched != 2) p.ksplit = small_ks, p.tail_begin = 0;
  } else if (p.epi_tma == 2 || p.epi_tma == 4) {
    // QKV: only the Q|K columns [0, 2D) are stored through the map; GEGLU: N/2 output columns
    const uint64_t dims[2] = {(uint64_t)(epi->store_mode == ST_QKV ? 2 * epi->qkv_d : p.n_out), (uint64_t)M};
    const uint64_t str[1] = {(uint64_t)epi->ldo * 2};
    const uint32_t box[2] = {32, 128};
    if ((rc = encode_map(&p.tmD, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, CU_TENSOR_MAP_SWIZZLE_64B, epi->out, 2, dims, str, box)))
      return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)M};
    const uint64_t str[1] = {(uint64_t)lda * 2};
    const uint32_t box[2] = {64, 128};
    if ((rc = encode_bf16_map(&p.tmA, A, 2, dims, str, box))) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)K, (uint64_t)N};
    const uint64_t str[1] = {(uint64_t)ldw * 2};
    const uint32_t box[2] = {64, (uint32_t)(two ? bn / 2 : bn)};
    if ((rc = encode_bf16_map(&p.tmB, W, 2, dims, str, box))) return rc;
  }
  return check_cuda(launch_gemm(p, bn, two, num_sms(), static_cast<cudaStream_t>(stream)), "gemm launch");
}

int vda_conv3x3_nhwc(const void* x, int ldx, const void* w, int n, int H, int W, int C, int Cout,
                     const vda_epilogue* epi, void* stream) {
  if (!x || !w || !epi || n <= 0 || H <= 0 || W <= 0) return set_error(-1, "vda_conv3x3_nhwc: bad arguments");
  if (C % 64 || ldx % 8 || ldx < C) return set_error(-1, "vda_conv3x3_nhwc: C=%d must be a multiple of 64 (ldx=%d)", C, ldx);
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = n * H * W, p.N = Cout, p.K = 9 * C;
  p.a_mode = A_CONV3X3;
  p.H = H, p.W = W, p.C = C;
  p.num_k_blocks = 9 * (C / 64);
  // spatial tile TH x TW = 128 pixels: pick the power-of-two TW with the fewest tiles
  long long best = -1;
  for (int l = 3; l <= 7; ++l) {
    const int tw = 1 << l, th = 128 >> l;
    const long long t = (long long)((W + tw - 1) / tw) * ((H + th - 1) / th);
    if (best < 0 || t < best) best = t, p.tw_log2 = l;
  }
  const int tw = 1 << p.tw_log2, th = 128 >> p.tw_log2;
  p.tiles_w = (W + tw - 1) / tw, p.tiles_h = (H + th - 1) / th;
  p.m_tiles = n * p.tiles_w * p.tiles_h;
  int rc = fill_epilogue(p, epi);
  if (rc) return rc;
  if (epi->store_mode != ST_PLAIN && epi->store_mode != ST_HEADDOT) return set_error(-1, "conv3x3: unsupported store mode");
  int bn;
  bool two;
  pick_tile(epi, Cout, p.m_tiles, &bn, &two);
  p.n_tiles = (Cout + bn - 1) / bn;
  p.epi_tma = pick_epi_tma(epi, true);
  if (p.epi_tma == 5 && Cout % 32) p.epi_tma = 0;   // whole 32-column boxes only
  p.stages = pick_stages(bn, two, epi, p.epi_tma);
  // halo mode (gemm.cu): one halo patch per channel block + tap-shifted descriptors instead of nine separate
  // (tap, block) A tiles.  1-CTA kernel: few output channels (BN <= 64; a stage = patch + its 9 W tiles).
  // CTA-pair kernel: BN = 128 (output_conv1), W tiles in their own ring; needs the bf16 box-store epilogue.
  {
    const char* hv = getenv("VDA_CONV_HALO");  // A/B knob: 0 disables
    p.halo = !(hv && atoi(hv) == 0) && !two && bn <= 64 && C % 64 == 0;
    if (p.halo) {
      const int halo_stage = 23 * 1024 + 9 * bn * 128;
      const int max_halo = bn <= 32 ? 3 : 2;
      const bool staged = p.epi_tma || (epi->res_f32 && epi->store_mode == ST_PLAIN && epi->act != ACT_GEGLU);
      p.stages = staged ? 2 : max_halo;  // the epilogue's 48 KB of boxes / residual ring sit behind the ring
      const int ring = (bn <= 32 ? 8 * 20480 : 8 * 24576) > max_halo * halo_stage ? (bn <= 32 ? 8 * 20480 : 8 * 24576) : max_halo * halo_stage;
      const int epi_need = p.epi_tma ? epi_box_area_bytes(p.epi_tma) : 48 * 1024;
      if (staged && p.stages * halo_stage + epi_need > ring) p.halo = 0, p.stages = pick_stages(bn, two, epi, p.epi_tma);
    }
    const char* h2 = getenv("VDA_CONV_HALO2");  // A/B knob: 0 off, 1 BN = 128 only, 2 (default) BN = 128 and 256
    const int h2v = h2 ? atoi(h2) : 2;
    const bool res_staged = !p.epi_tma && epi->res_f32 && epi->store_mode == ST_PLAIN && epi->act != ACT_GEGLU;
    if (two && C % 64 == 0 && (p.epi_tma == 2 || p.epi_tma == 5 || (p.epi_tma == 0 && !res_staged)) &&
        ((bn == 128 && h2v >= 1) || (bn == 256 && h2v >= 2))) {
      const int ring = bn >= 256 ? 6 * 32768 : 8 * 24576;            // Gemm2Cfg<BN>::kStages * kStageBytes
      const int w_bytes = (bn / 2) * 128;
      int nw = (ring - (p.epi_tma ? epi_box_area_bytes(p.epi_tma) : 0) - 2 * 23 * 1024) / w_bytes;
      if (nw > 16) nw = 16;
      if (nw >= 4) p.halo = 1, p.stages = nw;
    }
  }
  if (p.halo) {
    p.tw_log2 = 3;
    p.tiles_w = (W + 7) / 8, p.tiles_h = (H + 15) / 16;
    p.m_tiles = n * p.tiles_w * p.tiles_h;
    const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)n};
    const uint64_t str[3] = {(uint64_t)ldx * 2, (uint64_t)W * ldx * 2, (uint64_t)H * W * ldx * 2};
    const uint32_t box[4] = {64, 10, 18, 1};
    if ((rc = encode_bf16_map(&p.tmH, x, 4, dims, str, box))) return rc;
  }
  if (p.epi_tma == 2) {
    const uint64_t dims[4] = {(uint64_t)Cout, (uint64_t)W, (uint64_t)H, (uint64_t)n};
    const uint64_t str[3] = {(uint64_t)epi->ldo * 2, (uint64_t)W * epi->ldo * 2, (uint64_t)H * W * epi->ldo * 2};
    const uint32_t box[4] = {32, 1u << p.tw_log2, 128u >> p.tw_log2, 1};
    if ((rc = encode_map(&p.tmD, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, CU_TENSOR_MAP_SWIZZLE_64B, epi->out, 4, dims, str, box)))
      return rc;
  }
  {
    const uint64_t dims[4] = {(uint64_t)C, (uint64_t)W, (uint64_t)H, (uint64_t)n};
    const uint64_t str[3] = {(uint64_t)ldx * 2, (uint64_t)W * ldx * 2, (uint64_t)H * W * ldx * 2};
    const uint32_t box[4] = {64, 1u << p.tw_log2, 128u >> p.tw_log2, 1};
    if ((rc = encode_bf16_map(&p.tmA, x, 4, dims, str, box))) return rc;
  }
  {
    const uint64_t dims[2] = {(uint64_t)(9 * C), (uint64_t)Cout};
    const uint64_t str[1] = {(uint64_t)(9 * C) * 2};
    const uint32_t box[2] = {64, (uint32_t)(two ? bn / 2 : bn)};
    if ((rc = encode_bf16_map(&p.tmB, w, 2, dims, str, box))) return rc;
  }
  return check_cuda(launch_gemm(p, bn, two, num_sms(), static_cast<cudaStream_t>(stream)), "conv3x3 launch");
}

}  // extern "C"

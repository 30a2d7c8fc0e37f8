// C-ABI entry points that are not kernels themselves: error strings, launch
// accounting, workspace sizing and the tmx_prefill_attn dispatcher.
#include <cstdlib>
#include <mutex>
#include <string>
#include "tmx_attn.cuh"

namespace tmx {

static thread_local int g_launches = 0;
static std::mutex g_err_mu;
static std::string g_cuda_err;

void count_launch(int n) { g_launches += n; }
void reset_launch_count() { g_launches = 0; }

void set_cuda_error(cudaError_t e, const char* where) {
  std::lock_guard<std::mutex> lk(g_err_mu);
  g_cuda_err = std::string(cudaGetErrorName(e)) + ": " + cudaGetErrorString(e) + " at " + where;
}

int device_sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

int sm100_variant() {
  static int v = -1;
  if (v < 0) {
    const char* e = std::getenv("TMX_SM100_VARIANT");
    v = e ? (std::atoi(e) & 255) : 0;
  }
  return v;
}

int check_arch() {
  static int status = -1;
  if (status < 0) {
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) {
      set_cuda_error(cudaGetLastError(), "cudaDeviceGetAttribute");
      return TMX_ERR_CUDA;
    }
    status = (major == 10) ? TMX_OK : TMX_ERR_ARCH;
  }
  return status;
}

namespace {
size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct Plan {
  int kernel;      // TMX_KERNEL_MMA / TMX_KERNEL_TCGEN05
  int bm, bn;
  int stream;      // triangle layer: streaming tcgen05 kernel + mma kernel on the tail tiles only
  int stream_bm;   // tokens per streaming CTA
  int tail_tc;     // dense tail rows on the tcgen05 split-KV path (else the mma.sync kernel)
  int tail_bm;     // query-tile size of the tail path
  int n_slots, last_tiles, nsplit_max;
  size_t off_slots, off_part_o, off_part_ml, total;
};

Plan make_plan(int B, int total_T, int max_seqlen, int Hq, int Hkv, int si, int sl, int last, int mode,
               int kernel, bool allow_stream, bool allow_tail_tc) {
  Plan pl{};
  // AUTO: the tcgen05 kernel serves every tensor-bound case it can pair heads for (dense-causal
  // and unmasked layers).  Triangle layers are HBM-bound: their sink + window part goes to the
  // streaming tcgen05 kernel where it applies, the dense tail rows (and every other triangle
  // setting) to the mma.sync kernel with split-KV.
  const bool pairable = Hkv > 0 && Hq % Hkv == 0 && (Hq / Hkv) % 2 == 0;
  pl.kernel = kernel;
  if (kernel == TMX_KERNEL_AUTO)
    pl.kernel = (pairable && mode != TMX_MODE_TRIANGLE) ? TMX_KERNEL_TCGEN05 : TMX_KERNEL_MMA;
  pl.bn = 64;
  pl.bm = (mode == TMX_MODE_TRIANGLE) ? 64 : 128;
  if (pl.kernel == TMX_KERNEL_TCGEN05) { pl.bm = 128; pl.bn = sm100_block_n(sm100_variant()); }
  pl.stream = allow_stream && kernel == TMX_KERNEL_AUTO && mode == TMX_MODE_TRIANGLE &&
              stream_supported(Hq, Hkv, si, sl) && (last <= 1 || last_tiles_max(last, pl.bm) > 0);
  pl.stream_bm = pl.stream ? stream_block_m(Hq, Hkv) : 0;
  pl.n_slots = total_T / (pl.stream ? pl.stream_bm : pl.bm) + B;
  pl.last_tiles = 0; pl.nsplit_max = 0;
  pl.tail_bm = pl.bm;
  pl.tail_tc = pl.stream && allow_tail_tc && tail_tc_supported(Hq, Hkv, 1.0f) &&
               last_tiles_max(last, tail_tc_block_m(Hq, Hkv)) > 0;
  if (pl.tail_tc) {
    // one CTA = all query heads of a KV head x one KV chunk; aim at one wave of CTAs
    pl.tail_bm = tail_tc_block_m(Hq, Hkv);
    pl.last_tiles = ceil_div(last - 1, pl.tail_bm);      // tail tiles start at the first dense row
    int ns = device_sm_count() / (B * Hkv > 0 ? B * Hkv : 1);
    pl.nsplit_max = tail_nsplit(ceil_div(max_seqlen, 128), ns < 1 ? 1 : ns);
  } else if (pl.kernel == TMX_KERNEL_MMA && mode == TMX_MODE_TRIANGLE) {
    int ns = nsplit_max(max_seqlen, pl.bn);
    int lt = last_tiles_max(last, pl.bm);
    if (lt > 0 && ceil_div(max_seqlen, pl.bn) >= kSplitMinTiles) { pl.last_tiles = lt; pl.nsplit_max = ns; }
    else if (pl.stream) pl.last_tiles = lt;          // tail tiles without split workers
  }
  size_t off = 0;
  pl.off_slots = off; off = align_up(off + sizeof(SlotInfo) * (size_t)pl.n_slots, 256);
  size_t parts = (size_t)B * pl.last_tiles * pl.nsplit_max * Hq;
  pl.off_part_o = off; off = align_up(off + parts * pl.tail_bm * 128 * sizeof(float), 256);
  pl.off_part_ml = off; off = align_up(off + parts * pl.tail_bm * sizeof(float2), 256);
  pl.total = off;
  return pl;
}
}  // namespace
}  // namespace tmx

using namespace tmx;

extern "C" const char* tmx_strerror(int status) {
  switch (status) {
    case TMX_OK: return "ok";
    case TMX_ERR_BAD_ARG: return "bad argument (null pointer, negative size or unknown enum)";
    case TMX_ERR_BAD_SHAPE: return "unsupported shape (head_dim must be 128, Hq % Hkv == 0, strides >= H*D)";
    case TMX_ERR_ALIGNMENT: return "pointer or stride not 16-byte aligned";
    case TMX_ERR_WORKSPACE: return "workspace missing or smaller than tmx_attn_workspace_bytes()";
    case TMX_ERR_ARCH: return "device is not compute capability 10.x (B200, sm_100a)";
    case TMX_ERR_CUDA: return "CUDA call failed (see tmx_last_cuda_error)";
    case TMX_ERR_UNSUPPORTED: return "request not supported by this build";
    default: return "unknown status";
  }
}

extern "C" const char* tmx_last_cuda_error(void) {
  static thread_local std::string copy;
  std::lock_guard<std::mutex> lk(g_err_mu);
  copy = g_cuda_err;
  return copy.c_str();
}

extern "C" int tmx_version(void) { return 100; }
extern "C" int tmx_last_launch_count(void) { return g_launches; }

extern "C" size_t tmx_attn_workspace_bytes(int B, int total_T, int max_seqlen, int Hq, int Hkv,
                                           int D, int si, int sl, int last, int mode, int kernel) {
  (void)D;
  if (B <= 0 || total_T <= 0) return 256;
  // a bitmask (unknown here) disables the streaming plan: size for whichever plan is larger
  size_t best = 0;
  for (int v = 0; v < 3; ++v) {
    const size_t t = make_plan(B, total_T, max_seqlen, Hq, Hkv, si, sl, last, mode, kernel, v > 0, v > 1).total;
    best = t > best ? t : best;
  }
  return best;
}

extern "C" int tmx_prefill_attn(void* out, const void* q, const void* k, const void* v,
                                const int32_t* cu_seqlens, int B, int total_T, int max_seqlen,
                                int Hq, int Hkv, int D, int64_t q_stride_t, int64_t kv_stride_t,
                                int64_t o_stride_t, float scale, int si, int sl, int last,
                                int mode, const uint32_t* bitmask, int64_t bitmask_stride_words,
                                float* lse, void* workspace, size_t workspace_bytes, int kernel,
                                tmx_stream_t stream) {
  reset_launch_count();
  if (B < 0 || total_T < 0 || max_seqlen < 0 || Hq <= 0 || Hkv <= 0) return TMX_ERR_BAD_ARG;
  if (mode < TMX_MODE_DENSE_CAUSAL || mode > TMX_MODE_NONE) return TMX_ERR_BAD_ARG;
  if (kernel < TMX_KERNEL_AUTO || kernel > TMX_KERNEL_TCGEN05) return TMX_ERR_BAD_ARG;
  if (D != 128 || Hq % Hkv != 0) return TMX_ERR_BAD_SHAPE;
  if (q_stride_t < (int64_t)Hq * D || kv_stride_t < (int64_t)Hkv * D || o_stride_t < (int64_t)Hq * D)
    return TMX_ERR_BAD_SHAPE;
  if (B == 0 || total_T == 0) return TMX_OK;
  if (!out || !q || !k || !v || !cu_seqlens) return TMX_ERR_BAD_ARG;
  if (max_seqlen <= 0 || max_seqlen > total_T) return TMX_ERR_BAD_ARG;
  auto mis = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) != 0; };
  if (mis(out) || mis(q) || mis(k) || mis(v) || (q_stride_t % 8) || (kv_stride_t % 8) || (o_stride_t % 8))
    return TMX_ERR_ALIGNMENT;
  if (bitmask && (bitmask_stride_words * 32 < max_seqlen || bitmask_stride_words % 4)) return TMX_ERR_BAD_SHAPE;
  int st = check_arch();
  if (st != TMX_OK) return st;

  Plan pl = make_plan(B, total_T, max_seqlen, Hq, Hkv, si, sl, last, mode, kernel, bitmask == nullptr,
                      scale * 1.4426950408889634f > 0.f);
  if (!workspace || workspace_bytes < pl.total || mis(workspace)) return TMX_ERR_WORKSPACE;
  cudaStream_t s = (cudaStream_t)stream;
  uint8_t* ws = static_cast<uint8_t*>(workspace);

  AttnParams p{};
  p.q = (const __nv_bfloat16*)q; p.k = (const __nv_bfloat16*)k; p.v = (const __nv_bfloat16*)v;
  p.out = (__nv_bfloat16*)out; p.lse = lse; p.cu = cu_seqlens;
  p.slots = reinterpret_cast<SlotInfo*>(ws + pl.off_slots);
  p.bitmask = bitmask; p.bitmask_stride = bitmask_stride_words;
  p.part_o = reinterpret_cast<float*>(ws + pl.off_part_o);
  p.part_ml = reinterpret_cast<float2*>(ws + pl.off_part_ml);
  p.q_stride = q_stride_t; p.kv_stride = kv_stride_t; p.o_stride = o_stride_t;
  p.B = B; p.total_T = total_T; p.max_seqlen = max_seqlen; p.Hq = Hq; p.Hkv = Hkv;
  p.n_slots = pl.n_slots; p.last_tiles = pl.last_tiles; p.nsplit_max = pl.nsplit_max;
  p.si = si; p.sl = sl; p.last = last; p.mode = mode;
  p.scale_log2 = scale * 1.4426950408889634f;

  if (pl.stream) {
    // triangle layer = streaming kernel (rows before the first tail tile) + tail tiles
    // rows the streaming kernel leaves to the tail path start at a multiple of p.tail_bm: the
    // tcgen05 tail tiles begin exactly at the first dense row (1), the mma.sync ones are aligned
    p.tail_bm = pl.tail_tc ? 1 : pl.tail_bm;
    st = build_slots(reinterpret_cast<SlotInfo*>(ws + pl.off_slots), pl.n_slots, cu_seqlens, B,
                     max_seqlen, pl.stream_bm, pl.bn, si, sl, last, mode, s);
    if (st != TMX_OK) return st;
    // TMX_SM100_VARIANT bit 5: the one-CTA-per-tile streaming kernel (v1) instead of the persistent one
    st = (sm100_variant() & 32) ? launch_attn_stream(p, s) : launch_attn_stream2(p, s);
    if (st == TMX_ERR_UNSUPPORTED) st = launch_attn_stream(p, s);
    if (st != TMX_OK || last <= 1 || pl.last_tiles == 0) return st;
    if (pl.tail_tc) return launch_attn_tail_tc(p, s);
    p.tail_only = 1;
    p.n_slots = B * pl.last_tiles;
    return launch_attn_mma(p, pl.bm, s);
  }
  st = build_slots(reinterpret_cast<SlotInfo*>(ws + pl.off_slots), pl.n_slots, cu_seqlens, B,
                   max_seqlen, pl.bm, pl.bn, si, sl, last, mode, s);
  if (st != TMX_OK) return st;
  if (pl.kernel == TMX_KERNEL_TCGEN05) {
    const int var = sm100_variant();
    return sm100_use_v2(var) ? launch_attn_sm100(p, var & 15, s) : launch_attn_sm100_v3(p, var & 127, s);
  }
  return launch_attn_mma(p, pl.bm, s);
}

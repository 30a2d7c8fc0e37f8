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

// ---- per-device caches -----------------------------------------------------------------------
// One process may drive several GPUs (ops.py wraps every call in torch.cuda.device(...)): the SM
// count, the architecture check and "max dynamic shared memory already raised for kernel X" are
// properties of the CURRENT device, so they are kept per device ordinal under a mutex.
namespace {
constexpr int kMaxDevices = 64;
struct DeviceCache {
  int sms = 0;
  int arch_status = -1;
  int smem_attr[kAttrSlots] = {};
};
std::mutex g_dev_mu;
DeviceCache g_dev[kMaxDevices];
int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) dev = 0;
  return dev;
}
}  // namespace

int device_sm_count() {
  const int dev = current_device();
  std::lock_guard<std::mutex> lk(g_dev_mu);
  DeviceCache& c = g_dev[dev];
  if (c.sms == 0) {
    cudaDeviceGetAttribute(&c.sms, cudaDevAttrMultiProcessorCount, dev);
    if (c.sms <= 0) c.sms = 148;
  }
  return c.sms;
}

bool smem_attr_cached(int slot, int smem) {
  if (slot < 0 || slot >= kAttrSlots) return false;
  const int dev = current_device();
  std::lock_guard<std::mutex> lk(g_dev_mu);
  int& have = g_dev[dev].smem_attr[slot];
  if (have >= smem) return true;
  have = smem;
  return false;
}

int sm100_variant() {
#ifdef TMX_DEV
  static int v = -1;
  if (v < 0) {
    const char* e = std::getenv("TMX_SM100_VARIANT");
    v = e ? (std::atoi(e) & 255) : 0;
  }
  return v;
#else
  return 0;     // the product library ignores the environment
#endif
}

int check_arch() {
  const int dev = current_device();
  std::lock_guard<std::mutex> lk(g_dev_mu);
  DeviceCache& c = g_dev[dev];
  if (c.arch_status < 0) {
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) {
      set_cuda_error(cudaGetLastError(), "cudaDeviceGetAttribute");
      return TMX_ERR_CUDA;
    }
    c.arch_status = (major == 10) ? TMX_OK : TMX_ERR_ARCH;
  }
  return c.arch_status;
}

namespace {
size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct Plan {
  int kernel;      // TMX_KERNEL_MMA / TMX_KERNEL_TCGEN05: kernel of the block-index pass
  int bm, bn;
  int stream;      // triangle layer: streaming tcgen05 kernel for the rows before the dense tail
  int stream_bm;   // tokens per streaming CTA
  int tail_tc;     // dense tail rows on the tcgen05 split-KV path
  int tail_bm;     // query-tile size of the tail path
  int pair;        // block-index pass on CTA pairs (cta_group::2)
  int n_slots, last_tiles, nsplit_max;
  size_t off_slots, off_part_o, off_part_ml, total;
};

// How a call is served.  AUTO picks tcgen05 wherever two query heads share a KV head and the
// scale is positive (the tcgen05 kernels take the row max on raw scores):
//   dense-causal / unmasked   attn_sm100_pair_kernel (GQA ratio % 4 == 0) or attn_sm100_v3_kernel over the block index
//   triangle, sink < 16 and window <= 112 keys (the reference defaults)
//                             attn_stream3_kernel (HBM-bound pass) + split-KV tail on tcgen05
//   any other triangle        attn_sm100_v3_kernel over the block index (rows before the dense
//                             tail) + split-KV tail on tcgen05 while last - 1 <= kTailTcMaxRows
// mma.sync serves odd GQA ratios, non-positive scales and explicit kernel="mma" (the cross-check).
// `ctx`: the call carries context keys (chunked prefill / decode): block-index kernels only.
// `D`: head_dim.  The tcgen05 kernels are built for 128 (Qwen3); 64 is served by the mma.sync kernel alone.
Plan make_plan(int B, int total_T, int max_seqlen, int Hq, int Hkv, int si, int sl, int last, int mode,
               int kernel, bool has_bitmask, bool scale_pos, bool ctx, bool paged = false, int D = 128) {
  Plan pl{};
  const bool pairable = D == 128 && Hkv > 0 && Hq % Hkv == 0 && (Hq / Hkv) % 2 == 0;
  pl.kernel = kernel;
  // explicit request for the CTA-pair kernel (AUTO's choice between the two tcgen05 kernels: see below)
  pl.pair = kernel == TMX_KERNEL_TCGEN05_PAIR;
  if (pl.pair) pl.kernel = kernel = TMX_KERNEL_TCGEN05;
  if (kernel == TMX_KERNEL_AUTO) {
    pl.kernel = (pairable && scale_pos) ? TMX_KERNEL_TCGEN05 : TMX_KERNEL_MMA;
    // block-index pass: CTA pairs (cta_group::2) wherever four query heads share a KV head -- +6 % on the
    // 128K dense layer over the single-CTA kernel (profiles/r02_pair_kernel.md); paged K/V stay on the
    // single-CTA kernel (one TMA box per page)
    pl.pair = pl.kernel == TMX_KERNEL_TCGEN05 && (Hq / Hkv) % 4 == 0 && !paged;
  }
  const bool tri = mode == TMX_MODE_TRIANGLE;
  pl.bn = 64;
  pl.bm = tri ? 64 : 128;
  if (pl.kernel == TMX_KERNEL_TCGEN05) { pl.bm = 128; pl.bn = 128; }
  const bool tc_tri = kernel == TMX_KERNEL_AUTO && pl.kernel == TMX_KERNEL_TCGEN05 && tri && !has_bitmask && !ctx;
  pl.tail_tc = tc_tri && tail_tc_supported(Hq, Hkv, scale_pos ? 1.0f : 0.f) && last - 1 >= 1 &&
               last - 1 <= kTailTcMaxRows;
  pl.stream = tc_tri && stream_supported(Hq, Hkv, si, sl) && (last <= 1 || pl.tail_tc);
  pl.stream_bm = pl.stream ? stream_block_m(Hq, Hkv) : 0;
  pl.n_slots = total_T / (pl.stream ? pl.stream_bm : pl.bm) + B;
  pl.last_tiles = 0; pl.nsplit_max = 0;
  pl.tail_bm = pl.bm;
  if (pl.tail_tc) {
    // one CTA = all query heads of a KV head x one tail tile x one KV chunk; aim at one wave of CTAs
    pl.tail_bm = tail_tc_block_m(Hq, Hkv);
    pl.last_tiles = ceil_div(last - 1, pl.tail_bm);      // tail tiles start at the first dense row
    const long long ctas = (long long)B * Hkv * pl.last_tiles;
    int ns = (int)(device_sm_count() / (ctas > 0 ? ctas : 1));
    pl.nsplit_max = tail_nsplit(ceil_div(max_seqlen, 128), ns < 1 ? 1 : ns);
  } else if (pl.kernel == TMX_KERNEL_MMA && tri && !ctx) {
    int ns = nsplit_max(max_seqlen, pl.bn);
    int lt = last_tiles_max(last, pl.bm);
    if (lt > 0 && ceil_div(max_seqlen, pl.bn) >= kSplitMinTiles) { pl.last_tiles = lt; pl.nsplit_max = ns; }
  }
  size_t off = 0;
  pl.off_slots = off; off = align_up(off + sizeof(SlotInfo) * (size_t)pl.n_slots, 256);
  size_t parts = (size_t)B * pl.last_tiles * pl.nsplit_max * Hq;
  pl.off_part_o = off; off = align_up(off + parts * pl.tail_bm * 128 * sizeof(float), 256);
  pl.off_part_ml = off; off = align_up(off + parts * pl.tail_bm * sizeof(float2), 256);
  pl.total = off;
  return pl;
}

#ifdef TMX_DEV
long long* g_trace = nullptr;
#endif

// Shared by both entry points once the arguments are validated and AttnParams is filled.
int run_plan(const Plan& pl, AttnParams& p, uint8_t* ws, cudaStream_t s) {
#ifdef TMX_DEV
  p.trace = g_trace;
#endif
  p.slots = reinterpret_cast<SlotInfo*>(ws + pl.off_slots);
  p.part_o = reinterpret_cast<float*>(ws + pl.off_part_o);
  p.part_ml = reinterpret_cast<float2*>(ws + pl.off_part_ml);
  p.n_slots = pl.n_slots; p.last_tiles = pl.last_tiles; p.nsplit_max = pl.nsplit_max;
  SlotInfo* slots = reinterpret_cast<SlotInfo*>(ws + pl.off_slots);
  int st;
  if (pl.stream) {
    // triangle layer = streaming kernel (rows before the first dense row) + tail tiles, which
    // begin exactly at the first dense row (tail_bm = 1: no alignment)
    p.tail_bm = 1; p.tail_clip = 0;
    st = build_slots(slots, pl.n_slots, p, pl.stream_bm, pl.bn, s);
    if (st != TMX_OK) return st;
    st = launch_attn_stream3(p, s);
    if (st != TMX_OK || p.last <= 1 || !pl.tail_tc) return st;
    return launch_attn_tail_tc(p, s);
  }
  p.tail_clip = pl.tail_tc ? 1 : 0;
  st = build_slots(slots, pl.n_slots, p, pl.bm, pl.bn, s);
  if (st != TMX_OK) return st;
  if (pl.kernel == TMX_KERNEL_TCGEN05 && pl.pair) st = launch_attn_sm100_pair(p, s);
  else if (pl.kernel == TMX_KERNEL_TCGEN05) st = launch_attn_sm100_v3(p, sm100_variant() & 127, s);
  else st = launch_attn_mma(p, pl.bm, s);
  if (st != TMX_OK || !pl.tail_tc) return st;
  return launch_attn_tail_tc(p, s);
}
}  // namespace
}  // namespace tmx

using namespace tmx;

extern "C" const char* tmx_strerror(int status) {
  switch (status) {
    case TMX_OK: return "ok";
    case TMX_ERR_BAD_ARG: return "bad argument (null pointer, negative size or unknown enum)";
    case TMX_ERR_BAD_SHAPE: return "unsupported shape (head_dim must be 128 or 64, Hq % Hkv == 0, strides >= H*D)";
    case TMX_ERR_ALIGNMENT: return "pointer or stride not 16-byte aligned";
    case TMX_ERR_WORKSPACE: return "workspace missing or smaller than tmx_attn_workspace_bytes()";
    case TMX_ERR_ARCH: return "device is not compute capability 10.x (B200, sm_100a)";
    case TMX_ERR_CUDA: return "CUDA call failed (see tmx_last_cuda_error)";
    case TMX_ERR_UNSUPPORTED: return "request not supported by this build";
    case TMX_ERR_STATIC_SCRATCH:
      return "tmx_prefill_attn_static: the request needs more scratch than the 16 MiB static array "
             "(TMX_STATIC_SCRATCH_BYTES); call tmx_prefill_attn with a caller-provided workspace";
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

#ifdef TMX_DEV
// Developer library only: device buffer of 4 x 96 x 12 int64 that the streaming kernel fills with
// clock64 stamps of its pipeline events (tools/stream_trace.py); null switches the trace off.
extern "C" void tmx_dev_set_trace(long long* buf) { g_trace = buf; }
#endif

extern "C" size_t tmx_attn_workspace_bytes(int B, int total_T, int max_seqlen, int Hq, int Hkv,
                                           int D, int si, int sl, int last, int mode, int kernel) {
  if (B <= 0 || total_T <= 0) return 256;
  // a bitmask or the sign of the scale (unknown here) change the plan: size for the largest one
  size_t best = 0;
  for (int v = 0; v < 4; ++v) {
    const size_t t = make_plan(B, total_T, max_seqlen, Hq, Hkv, si, sl, last, mode, kernel, (v & 1) != 0,
                               (v & 2) != 0, false, false, D).total;
    best = t > best ? t : best;
  }
  return best;
}

namespace {
int check_common(int B, int total_T, int max_seqlen, int Hq, int Hkv, int D, int64_t q_stride_t,
                 int64_t kv_stride_t, int64_t o_stride_t, int mode, int kernel) {
  if (B < 0 || total_T < 0 || max_seqlen < 0 || Hq <= 0 || Hkv <= 0) return TMX_ERR_BAD_ARG;
  if (mode < TMX_MODE_DENSE_CAUSAL || mode > TMX_MODE_NONE) return TMX_ERR_BAD_ARG;
  if (kernel < TMX_KERNEL_AUTO || kernel > TMX_KERNEL_TCGEN05_PAIR) return TMX_ERR_BAD_ARG;
  if ((D != 128 && D != 64) || Hq % Hkv != 0) return TMX_ERR_BAD_SHAPE;
  if (D != 128 && kernel >= TMX_KERNEL_TCGEN05) return TMX_ERR_UNSUPPORTED;   // tcgen05 kernels: head_dim 128 only
  if (q_stride_t < (int64_t)Hq * D || kv_stride_t < (int64_t)Hkv * D || o_stride_t < (int64_t)Hq * D)
    return TMX_ERR_BAD_SHAPE;
  return TMX_OK;
}
bool misaligned(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) != 0; }
}  // namespace

extern "C" int tmx_prefill_attn(void* out, const void* q, const void* k, const void* v,
                                const int32_t* cu_seqlens, int B, int total_T, int max_seqlen,
                                int Hq, int Hkv, int D, int64_t q_stride_t, int64_t kv_stride_t,
                                int64_t o_stride_t, float scale, int si, int sl, int last,
                                int mode, const uint32_t* bitmask, int64_t bitmask_stride_words,
                                float* lse, void* workspace, size_t workspace_bytes, int kernel,
                                tmx_stream_t stream) {
  reset_launch_count();
  int st = check_common(B, total_T, max_seqlen, Hq, Hkv, D, q_stride_t, kv_stride_t, o_stride_t, mode, kernel);
  if (st != TMX_OK) return st;
  if (B == 0 || total_T == 0) return TMX_OK;
  if (!out || !q || !k || !v || !cu_seqlens) return TMX_ERR_BAD_ARG;
  if (max_seqlen <= 0 || max_seqlen > total_T) return TMX_ERR_BAD_ARG;
  if (misaligned(out) || misaligned(q) || misaligned(k) || misaligned(v) || (q_stride_t % 8) ||
      (kv_stride_t % 8) || (o_stride_t % 8))
    return TMX_ERR_ALIGNMENT;
  if (bitmask && (bitmask_stride_words * 32 < max_seqlen || bitmask_stride_words % 4)) return TMX_ERR_BAD_SHAPE;
  st = check_arch();
  if (st != TMX_OK) return st;

  const float scale_log2 = scale * 1.4426950408889634f;
  const bool scale_pos = scale_log2 > 0.f;
  // an explicit tcgen05 request cannot be served with a non-positive scale; AUTO falls back to mma.sync
  if (kernel >= TMX_KERNEL_TCGEN05 && !scale_pos) return TMX_ERR_UNSUPPORTED;
  Plan pl = make_plan(B, total_T, max_seqlen, Hq, Hkv, si, sl, last, mode, kernel, bitmask != nullptr,
                      scale_pos, false, false, D);
  if (!workspace || workspace_bytes < pl.total || misaligned(workspace)) return TMX_ERR_WORKSPACE;

  AttnParams p{};
  p.q = (const __nv_bfloat16*)q; p.k = (const __nv_bfloat16*)k; p.v = (const __nv_bfloat16*)v;
  p.out = (__nv_bfloat16*)out; p.lse = lse; p.cu = cu_seqlens;
  p.bitmask = bitmask; p.bitmask_stride = bitmask_stride_words;
  p.q_stride = q_stride_t; p.kv_stride = kv_stride_t; p.o_stride = o_stride_t;
  p.B = B; p.total_T = total_T; p.total_kv = total_T; p.max_seqlen = max_seqlen; p.Hq = Hq; p.Hkv = Hkv; p.D = D;
  p.si = si; p.sl = sl; p.last = last; p.mode = mode;
  p.scale_log2 = scale_log2;
  return run_plan(pl, p, static_cast<uint8_t*>(workspace), (cudaStream_t)stream);
}

extern "C" int tmx_chunked_prefill_attn(void* out, const void* q, const int32_t* cu_seqlens_q,
                                        const int32_t* seq_lens, int B, int total_Tq, int max_query_len,
                                        int Hq, int Hkv, int D, int64_t q_stride_t, int64_t o_stride_t,
                                        const void* k, const void* v, int64_t kv_stride_t, int64_t kv_rows,
                                        const int32_t* cu_seqlens_k, const int32_t* block_table,
                                        int64_t block_table_stride, int page_size, float scale, int si,
                                        int sl, int last, int mode, float* lse, void* workspace,
                                        size_t workspace_bytes, int kernel, tmx_stream_t stream) {
  reset_launch_count();
  int st = check_common(B, total_Tq, max_query_len, Hq, Hkv, D, q_stride_t, kv_stride_t, o_stride_t, mode, kernel);
  if (st != TMX_OK) return st;
  if (mode == TMX_MODE_NONE) return TMX_ERR_BAD_ARG;            // causal layers only
  if (B == 0 || total_Tq == 0) return TMX_OK;
  if (!out || !q || !k || !v || !cu_seqlens_q || !seq_lens || kv_rows <= 0) return TMX_ERR_BAD_ARG;
  if ((block_table == nullptr) == (cu_seqlens_k == nullptr)) return TMX_ERR_BAD_ARG;   // exactly one K/V addressing
  if (block_table && (page_size < 16 || page_size % 16 != 0 || block_table_stride <= 0)) return TMX_ERR_BAD_ARG;
  if (block_table && page_size < 128 && 128 % page_size != 0) return TMX_ERR_BAD_SHAPE;
  if (block_table && page_size > 128 && page_size % 128 != 0) return TMX_ERR_BAD_SHAPE;
  if (max_query_len <= 0 || max_query_len > total_Tq) return TMX_ERR_BAD_ARG;
  if (misaligned(out) || misaligned(q) || misaligned(k) || misaligned(v) || (q_stride_t % 8) ||
      (kv_stride_t % 8) || (o_stride_t % 8))
    return TMX_ERR_ALIGNMENT;
  st = check_arch();
  if (st != TMX_OK) return st;
  const float scale_log2 = scale * 1.4426950408889634f;
  const bool scale_pos = scale_log2 > 0.f;
  if (kernel >= TMX_KERNEL_TCGEN05 && !scale_pos) return TMX_ERR_UNSUPPORTED;
  Plan pl = make_plan(B, total_Tq, max_query_len, Hq, Hkv, si, sl, last, mode, kernel, false, scale_pos, true,
                      block_table != nullptr, D);
  if (!workspace || workspace_bytes < pl.total || misaligned(workspace)) return TMX_ERR_WORKSPACE;

  AttnParams p{};
  p.q = (const __nv_bfloat16*)q; p.k = (const __nv_bfloat16*)k; p.v = (const __nv_bfloat16*)v;
  p.out = (__nv_bfloat16*)out; p.lse = lse; p.cu = cu_seqlens_q;
  p.kv_lens = seq_lens; p.cu_k = cu_seqlens_k; p.block_table = block_table;
  p.bt_stride = block_table_stride; p.page_size = page_size;
  p.q_stride = q_stride_t; p.kv_stride = kv_stride_t; p.o_stride = o_stride_t;
  p.B = B; p.total_T = total_Tq; p.total_kv = kv_rows; p.max_seqlen = max_query_len; p.Hq = Hq; p.Hkv = Hkv; p.D = D;
  p.si = si; p.sl = sl; p.last = last; p.mode = mode;
  p.scale_log2 = scale_log2;
  return run_plan(pl, p, static_cast<uint8_t*>(workspace), (cudaStream_t)stream);
}

// ---- allocation-free variant of tmx_prefill_attn ----------------------------------------------------------
// No counterpart in the reference (include/tmx.h).  The scratch the
// attention kernels need (block index, split-KV partials) lives in this file-scope device array instead of a caller
// workspace.  In host code `g_static_scratch` is only a stand-in for the symbol: its DEVICE address is taken once per
// device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels get.
static __device__ __align__(256) unsigned char g_static_scratch[TMX_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kScratchMaxDevices = 64;
std::mutex g_scratch_mu;
void* g_scratch_addr[kScratchMaxDevices] = {};

int static_scratch_address(void** addr) {
  int dev = -1;
  TMX_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kScratchMaxDevices) return TMX_ERR_UNSUPPORTED;
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (g_scratch_addr[dev] == nullptr) {
    void* ptr = nullptr;
    TMX_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_static_scratch));      // `dev` is the current device
    if (ptr == nullptr) return TMX_ERR_CUDA;
    g_scratch_addr[dev] = ptr;
  }
  *addr = g_scratch_addr[dev];
  return TMX_OK;
}
}  // namespace

extern "C" int tmx_prefill_attn_static(void* out, const void* q, const void* k, const void* v,
                                       const int32_t* cu_seqlens, int B, int total_T, int max_seqlen,
                                       int Hq, int Hkv, int D, int64_t q_stride_t, int64_t kv_stride_t,
                                       int64_t o_stride_t, float scale, int si, int sl, int last,
                                       int mode, const uint32_t* bitmask, int64_t bitmask_stride_words,
                                       float* lse, int kernel, tmx_stream_t stream) {
  reset_launch_count();
  int st = check_common(B, total_T, max_seqlen, Hq, Hkv, D, q_stride_t, kv_stride_t, o_stride_t, mode, kernel);
  if (st != TMX_OK) return st;
  if (B == 0 || total_T == 0) return TMX_OK;
  if (max_seqlen <= 0 || max_seqlen > total_T) return TMX_ERR_BAD_ARG;
  // the same plan tmx_prefill_attn will make: its scratch must fit the static array
  const bool scale_pos = scale * 1.4426950408889634f > 0.f;
  const Plan pl = make_plan(B, total_T, max_seqlen, Hq, Hkv, si, sl, last, mode, kernel, bitmask != nullptr,
                            scale_pos, false, false, D);
  if (pl.total > (size_t)TMX_STATIC_SCRATCH_BYTES) return TMX_ERR_STATIC_SCRATCH;
  st = check_arch();
  if (st != TMX_OK) return st;
  void* scratch = nullptr;
  st = static_scratch_address(&scratch);
  if (st != TMX_OK) return st;
  return tmx_prefill_attn(out, q, k, v, cu_seqlens, B, total_T, max_seqlen, Hq, Hkv, D, q_stride_t, kv_stride_t,
                          o_stride_t, scale, si, sl, last, mode, bitmask, bitmask_stride_words, lse, scratch,
                          (size_t)TMX_STATIC_SCRATCH_BYTES, kernel, stream);
}

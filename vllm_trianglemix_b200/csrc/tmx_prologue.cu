// The step in front of attention, one HBM pass:
//   per-head q/k RMSNorm -> NeoX rotary embedding -> (optional) K/V write into the paged cache.
// Replaces, with the reference's rounding order (bf16 after every torch op of forward_native):
//   vllm/model_executor/models/qwen3.py:166-174                      (q_norm, k_norm, rotary_emb)
//   vllm/model_executor/layers/layernorm.py:170-206                  (RMSNorm.forward_native)
//   vllm/model_executor/layers/rotary_embedding/base.py:66-95        (forward_native, NeoX)
//   vllm/v1/attention/backends/flash_attn.py:498-507                 (reshape_and_cache_flash)
// HBM-bound, no reuse: q/k are read once and written once, v is read once and written once into
// the cache, the normed+rotated k goes to both the contiguous tensor the attention kernel reads
// and the cache.  One WARP per token (grid-stride): position, slot and the token's cos/sin row are
// fetched once and reused by all of its heads.  A head row (128 bf16 = 256 B) is served by a
// QUARTER-WARP: lane l8 holds elements [8*l8, 8*l8+8) and [64+8*l8, 64+8*l8+8) -- two 16-byte
// accesses, each a fully used 128-byte line per quarter-warp -- so both members of every rotary
// pair (e, e+64) sit in the same lane and the only exchange is a 3-step sum-of-squares butterfly.
// A warp covers 4 heads per pass and keeps two passes of loads in flight.  (First version: a
// half-warp per head with 8-byte accesses and a per-row decode ran at 70 % issue utilisation,
// 390 instructions per 4 rows -- profiles/r01_prologue_experiments.md.)
#include <cstdlib>
#include "tmx_common.cuh"

namespace tmx {
namespace {

constexpr int kD = 128;
constexpr int kThreads = 256;
constexpr int kWarpsPerCta = kThreads / 32;
constexpr int kUnroll = 2;
constexpr int kMinCtasPerSm = 3;        // <= 80 registers, no spills: 768 threads per SM, 64 B in flight each

struct PrologueParams {
  const __nv_bfloat16 *q, *k, *v;
  __nv_bfloat16 *q_out, *k_out;
  const __nv_bfloat16 *q_w, *k_w;
  const long long* pos;
  const __nv_bfloat16* cache;       // [max_pos, 128]: cos | sin
  __nv_bfloat16 *k_cache, *v_cache;
  const long long* slots;
  long long q_st, k_st, v_st, qo_st, ko_st;
  long long blk_st, tok_st, head_st;  // cache strides (elements)
  int T, Hq, Hkv, block_size, max_pos;
  int norm_rope;                      // 0: cache write only
  int bs_shift;                       // log2(block_size) when it is a power of two, else -1
  float eps;
};

__device__ __forceinline__ uint4 ld16(const __nv_bfloat16* p) {
  return __ldg(reinterpret_cast<const uint4*>(p));
}
__device__ __forceinline__ uint4 ld16_stream(const __nv_bfloat16* p) {  // read-once data: keep it out of L1
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void st16(__nv_bfloat16* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }

// ---- arithmetic in the reference's rounding order without the conversion (XU) pipe --------
// The reference rounds to bf16 after every torch op.  For two bf16 operands the fp32 product is
// exact, and the fp32 sum/difference is either exact or differs from the exact value by less than
// the bf16 rounding could see, so "fp32 op, then round to bf16" == the single-rounded native
// bf16x2 instruction (mul.rn.bf16x2 / add.rn.bf16x2 / sub.rn.bf16x2).  Only x * rsqrt(var) has an
// fp32 operand; that one is an FMUL followed by the packed cvt.rn.bf16x2.f32.
__device__ __forceinline__ float lo_f(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float hi_f(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_rn(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ uint32_t mul2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t add2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("add.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t sub2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("sub.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

struct Row {
  uint4 a, b;  // elements [8*l8, 8*l8+8) and [64+8*l8, 64+8*l8+8)
};

// RMSNorm + NeoX RoPE of one head row held across a quarter-warp (qmask names its 8 lanes).
// wp -> this lane's 8 + 8 weight elements (L1-resident; not kept in registers across rows).
__device__ __forceinline__ Row norm_rope_row(const Row& x, const __nv_bfloat16* wp, const uint4& cosv,
                                             const uint4& sinv, float eps, unsigned qmask) {
  const uint32_t xw[8] = {x.a.x, x.a.y, x.a.z, x.a.w, x.b.x, x.b.y, x.b.z, x.b.w};
  float s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int e = 0; e < 8; ++e) {                         // bf16 squares are exact in fp32
    const float lo = lo_f(xw[e]), hi = hi_f(xw[e]);
    s0 = fmaf(lo, lo, s0);
    s1 = fmaf(hi, hi, s1);
  }
  float ss = s0 + s1;
#pragma unroll
  for (int o = 4; o >= 1; o >>= 1) ss += __shfl_xor_sync(qmask, ss, o);
  // torch.rsqrt on the CPU is 1 / sqrt(x), both correctly rounded; so are these
  const float rs = __frcp_rn(__fsqrt_rn(ss * (1.0f / kD) + eps));
  const uint4 wa = ld16(wp), wb = ld16(wp + 64);
  const uint32_t ww[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
  const uint32_t c[4] = {cosv.x, cosv.y, cosv.z, cosv.w}, sn[4] = {sinv.x, sinv.y, sinv.z, sinv.w};
  uint32_t o1[4], o2[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {                         // x1 = words 0..3, x2 = words 4..7
    // (x * rsqrt).to(bf16) * weight
    const uint32_t n1 = mul2(pack_rn(lo_f(xw[e]) * rs, hi_f(xw[e]) * rs), ww[e]);
    const uint32_t n2 = mul2(pack_rn(lo_f(xw[4 + e]) * rs, hi_f(xw[4 + e]) * rs), ww[4 + e]);
    o1[e] = sub2(mul2(n1, c[e]), mul2(n2, sn[e]));      // x1*cos - x2*sin, each op rounded
    o2[e] = add2(mul2(n2, c[e]), mul2(n1, sn[e]));      // x2*cos + x1*sin
  }
  Row r;
  r.a = make_uint4(o1[0], o1[1], o1[2], o1[3]);
  r.b = make_uint4(o2[0], o2[1], o2[2], o2[3]);
  return r;
}

// One token per warp iteration.  The token's head rows form one list -- q heads, k heads, then
// (when the cache is written) v heads -- walked 4 rows per pass (one per quarter-warp), kUnroll
// passes in flight, so a rank that holds 4 + 1 + 1 rows per token (Qwen3-8B over 8 GPUs) issues
// them in one round trip.  The body is straight-line: rows past the end re-read the last row and
// only their stores are predicated off; v rows run through the norm arithmetic and keep the raw
// value (a select) instead of branching around it.
template <bool kNorm, int kMinCtas>
__global__ void __launch_bounds__(kThreads, kMinCtas) prologue_kernel(const PrologueParams p) {
  const int lane = threadIdx.x & 31;
  const int quarter = lane >> 3;
  const int e0 = 8 * (lane & 7);
  const unsigned qmask = 0xffu << (8 * quarter);
  const long long warp0 = (long long)blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * kWarpsPerCta;
  const int hqk = p.Hq + p.Hkv;
  const int rows = hqk + (p.slots ? p.Hkv : 0);

  for (long long t = warp0; t < p.T; t += nwarps) {
    long long slot = -1;
    if (p.slots) slot = p.slots[t];
    long long coff = 0;
    if (slot >= 0) {   // padding tokens carry slot -1 (cache_kernels.cu:292-297)
      const long long blk = p.bs_shift >= 0 ? (slot >> p.bs_shift) : slot / p.block_size;
      coff = blk * p.blk_st + (slot - blk * p.block_size) * p.tok_st + e0;
    }
    uint4 cosv = uint4{}, sinv = uint4{};
    if (kNorm) {
      const __nv_bfloat16* cs = p.cache + p.pos[t] * kD + e0;
      cosv = ld16(cs);
      sinv = ld16(cs + 64);
    }
    const __nv_bfloat16* qrow = p.q + t * p.q_st + e0;
    const __nv_bfloat16* krow = p.k + t * p.k_st + e0 - (long long)p.Hq * kD;          // indexed by list row j
    const __nv_bfloat16* vrow = p.v + t * p.v_st + e0 - (long long)hqk * kD;
    __nv_bfloat16* qdst = p.q_out + t * p.qo_st + e0;
    __nv_bfloat16* kdst = p.k_out + t * p.ko_st + e0 - (long long)p.Hq * kD;

#pragma unroll 1
    for (int j0 = 0; j0 < rows; j0 += 4 * kUnroll) {
      Row x[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int j = min(j0 + 4 * u + quarter, rows - 1);
        const __nv_bfloat16* src = (j < p.Hq ? qrow : j < hqk ? krow : vrow) + j * kD;
        x[u].a = ld16_stream(src);
        x[u].b = ld16_stream(src + 64);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int j = j0 + 4 * u + quarter;
        const bool is_q = j < p.Hq, is_v = j >= hqk;
        Row y = x[u];
        if (kNorm) {
          const Row n = norm_rope_row(x[u], (is_q ? p.q_w : p.k_w) + e0, cosv, sinv, p.eps, qmask);
          if (!is_v) y = n;
          if (j < hqk) {
            __nv_bfloat16* dst = (is_q ? qdst : kdst) + j * kD;
            st16(dst, y.a);
            st16(dst + 64, y.b);
          }
        }
        if (slot >= 0 && !is_q && j < rows) {
          __nv_bfloat16* dst = (is_v ? p.v_cache + (long long)(j - hqk) * p.head_st
                                     : p.k_cache + (long long)(j - p.Hq) * p.head_st) + coff;
          st16(dst, y.a);
          st16(dst + 64, y.b);
        }
      }
    }
  }
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

int launch(const PrologueParams& p, cudaStream_t s) {
  reset_launch_count();
  if (p.T == 0) return TMX_OK;
  int st = check_arch();
  if (st != TMX_OK) return st;
  static int ctas_per_sm = 0;
  static void (*kernel)(const PrologueParams) = nullptr;
  static void (*copy_kernel)(const PrologueParams) = nullptr;
  if (ctas_per_sm == 0) {
#ifdef TMX_DEV
    const char* e = std::getenv("TMX_PROLOGUE_VARIANT");   // developer switch: 1 = 4 CTAs/SM build (64 regs, small spill)
    const bool v1 = e && std::atoi(e) == 1;
#else
    const bool v1 = false;                                 // the product library reads no environment variable
#endif
    kernel = v1 ? prologue_kernel<true, 4> : prologue_kernel<true, kMinCtasPerSm>;
    copy_kernel = prologue_kernel<false, kMinCtasPerSm>;
    TMX_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, kernel, kThreads, 0));
    if (ctas_per_sm <= 0) ctas_per_sm = 1;
  }
  PrologueParams pp = p;
  pp.bs_shift = -1;
  for (int b = 0; b < 31; ++b)
    if (p.block_size == (1 << b)) pp.bs_shift = b;
  const long long want = ((long long)p.T + kWarpsPerCta - 1) / kWarpsPerCta;
  const long long cap = (long long)device_sm_count() * ctas_per_sm * 2;   // 2 waves of resident CTAs
  const unsigned grid = (unsigned)(want < cap ? want : cap);
  (p.norm_rope ? kernel : copy_kernel)<<<grid, kThreads, 0, s>>>(pp);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

}  // namespace
}  // namespace tmx

using namespace tmx;

extern "C" int tmx_qk_norm_rope(void* q_out, void* k_out, const void* q, const void* k,
                                const void* q_weight, const void* k_weight, const int64_t* positions,
                                const void* cos_sin_cache, int max_pos, int T, int Hq, int Hkv, int D,
                                int64_t q_stride_t, int64_t k_stride_t, int64_t qo_stride_t,
                                int64_t ko_stride_t, float eps, const void* v, int64_t v_stride_t,
                                void* k_cache, void* v_cache, const int64_t* slot_mapping,
                                int64_t cache_block_stride, int64_t cache_token_stride,
                                int64_t cache_head_stride, int block_size, tmx_stream_t stream) {
  if (T < 0 || Hq < 0 || Hkv < 0 || max_pos < 0) return TMX_ERR_BAD_ARG;
  if (D != kD) return TMX_ERR_BAD_SHAPE;
  if (T == 0 || Hq + Hkv == 0) return TMX_OK;
  if (!positions || !cos_sin_cache || max_pos == 0) return TMX_ERR_BAD_ARG;
  if (Hq > 0 && (!q || !q_out || !q_weight)) return TMX_ERR_BAD_ARG;
  if (Hkv > 0 && (!k || !k_out || !k_weight)) return TMX_ERR_BAD_ARG;
  const bool cache = k_cache != nullptr || v_cache != nullptr;
  if (cache && (!k_cache || !v_cache || !v || !slot_mapping || block_size <= 0 || Hkv == 0)) return TMX_ERR_BAD_ARG;
  if (q_stride_t < (int64_t)Hq * D || qo_stride_t < (int64_t)Hq * D || k_stride_t < (int64_t)Hkv * D ||
      ko_stride_t < (int64_t)Hkv * D)
    return TMX_ERR_BAD_SHAPE;
  if (cache && (v_stride_t < (int64_t)Hkv * D || cache_head_stride < D)) return TMX_ERR_BAD_SHAPE;
  const int64_t strides[] = {q_stride_t, k_stride_t, qo_stride_t, ko_stride_t,
                             cache ? v_stride_t : 0, cache ? cache_block_stride : 0,
                             cache ? cache_token_stride : 0, cache ? cache_head_stride : 0};
  for (int64_t sdt : strides)
    if (sdt % 8 != 0) return TMX_ERR_ALIGNMENT;
  const void* ptrs[] = {q_out, k_out, q, k, q_weight, k_weight, cos_sin_cache, v, k_cache, v_cache};
  for (const void* ptr : ptrs)
    if (!aligned16(ptr)) return TMX_ERR_ALIGNMENT;
  PrologueParams p{};
  p.q = (const __nv_bfloat16*)q; p.k = (const __nv_bfloat16*)k; p.v = (const __nv_bfloat16*)v;
  p.q_out = (__nv_bfloat16*)q_out; p.k_out = (__nv_bfloat16*)k_out;
  p.q_w = (const __nv_bfloat16*)q_weight; p.k_w = (const __nv_bfloat16*)k_weight;
  p.pos = (const long long*)positions; p.cache = (const __nv_bfloat16*)cos_sin_cache;
  p.k_cache = (__nv_bfloat16*)k_cache; p.v_cache = (__nv_bfloat16*)v_cache;
  p.slots = cache ? (const long long*)slot_mapping : nullptr;
  p.q_st = q_stride_t; p.k_st = k_stride_t; p.v_st = v_stride_t; p.qo_st = qo_stride_t; p.ko_st = ko_stride_t;
  p.blk_st = cache_block_stride; p.tok_st = cache_token_stride; p.head_st = cache_head_stride;
  p.T = T; p.Hq = Hq; p.Hkv = Hkv;
  p.block_size = block_size; p.max_pos = max_pos; p.norm_rope = 1; p.eps = eps;
  return launch(p, (cudaStream_t)stream);
}

extern "C" int tmx_reshape_and_cache_flash(const void* key, const void* value, void* k_cache, void* v_cache,
                                           const int64_t* slot_mapping, int T, int Hkv, int D,
                                           int64_t k_stride_t, int64_t v_stride_t,
                                           int64_t cache_block_stride, int64_t cache_token_stride,
                                           int64_t cache_head_stride, int block_size, tmx_stream_t stream) {
  if (T < 0 || Hkv < 0) return TMX_ERR_BAD_ARG;
  if (D != kD) return TMX_ERR_BAD_SHAPE;
  if (T == 0 || Hkv == 0) return TMX_OK;
  if (!key || !value || !k_cache || !v_cache || !slot_mapping || block_size <= 0) return TMX_ERR_BAD_ARG;
  if (k_stride_t < (int64_t)Hkv * D || v_stride_t < (int64_t)Hkv * D || cache_head_stride < D) return TMX_ERR_BAD_SHAPE;
  const int64_t strides[] = {k_stride_t, v_stride_t, cache_block_stride, cache_token_stride, cache_head_stride};
  for (int64_t sdt : strides)
    if (sdt % 8 != 0) return TMX_ERR_ALIGNMENT;
  if (!aligned16(key) || !aligned16(value) || !aligned16(k_cache) || !aligned16(v_cache)) return TMX_ERR_ALIGNMENT;
  PrologueParams p{};
  p.k = (const __nv_bfloat16*)key; p.v = (const __nv_bfloat16*)value;
  p.k_cache = (__nv_bfloat16*)k_cache; p.v_cache = (__nv_bfloat16*)v_cache;
  p.slots = (const long long*)slot_mapping;
  p.k_st = k_stride_t; p.v_st = v_stride_t;
  p.blk_st = cache_block_stride; p.tok_st = cache_token_stride; p.head_st = cache_head_stride;
  p.T = T; p.Hq = 0; p.Hkv = Hkv;
  p.block_size = block_size; p.norm_rope = 0;
  return launch(p, (cudaStream_t)stream);
}

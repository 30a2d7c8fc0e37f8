// The step in front of attention, one HBM pass:
//   per-head q/k RMSNorm -> NeoX rotary embedding -> (optional) K/V write into the paged cache.
// Replaces, with the reference's rounding order (bf16 after every torch op of forward_native):
//   vllm/model_executor/models/qwen3.py:166-174                      (q_norm, k_norm, rotary_emb)
//   vllm/model_executor/layers/layernorm.py:170-206                  (RMSNorm.forward_native)
//   vllm/model_executor/layers/rotary_embedding/base.py:66-95        (forward_native, NeoX)
//   vllm/v1/attention/backends/flash_attn.py:498-507                 (reshape_and_cache_flash)
// HBM-bound, no reuse: q/k are read once and written once, v is read once and written once into
// the cache, the normed+rotated k goes to both the contiguous tensor the attention kernel reads
// and the cache.  Work unit = one (token, head) row of 128 bf16 = 256 B, served by a HALF-WARP:
// lane l holds elements [4l, 4l+4) and [64+4l, 64+4l+4) (two 8-byte accesses, each a fully used
// 128-byte line per half-warp), so both members of every rotary pair (e, e+64) sit in the same
// lane and the only exchange is the 4-step sum-of-squares butterfly.  Two units are in flight
// per half-warp per iteration (loads of both issued before either is used) and the grid is a
// multiple of the SM count.
#include "tmx_common.cuh"

namespace tmx {
namespace {

constexpr int kD = 128;
constexpr int kThreads = 256;
constexpr int kHalfWarpsPerCta = kThreads / 16;
constexpr int kUnroll = 2;

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
  long long units;                    // T * units_per_token
  int T, Hq, Hkv, upt, block_size, max_pos;
  int norm_rope;                      // 0: cache write only
  float eps;
};

struct Row {
  uint2 a, b;  // elements [4l,4l+4) and [64+4l, 64+4l+4)
};

__device__ __forceinline__ uint2 ld8(const __nv_bfloat16* p) {
  return __ldg(reinterpret_cast<const uint2*>(p));
}
__device__ __forceinline__ uint2 ld8_stream(const __nv_bfloat16* p) {  // read-once data: keep it out of L1
  uint2 r;
  asm volatile("ld.global.L1::no_allocate.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, uint2 v) { *reinterpret_cast<uint2*>(p) = v; }

__device__ __forceinline__ void unpack4(uint2 r, float* f) {
  f[0] = __uint_as_float(r.x << 16);
  f[1] = __uint_as_float(r.x & 0xffff0000u);
  f[2] = __uint_as_float(r.y << 16);
  f[3] = __uint_as_float(r.y & 0xffff0000u);
}
// round-to-nearest-even to bf16 and back: "the torch op returned a bf16 tensor"
__device__ __forceinline__ float rbf(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
__device__ __forceinline__ uint2 pack4(const float* f) {  // f already bf16-representable
  uint2 r;
  r.x = (__float_as_uint(f[0]) >> 16) | (__float_as_uint(f[1]) & 0xffff0000u);
  r.y = (__float_as_uint(f[2]) >> 16) | (__float_as_uint(f[3]) & 0xffff0000u);
  return r;
}

struct Unit {
  long long t;
  int h;      // head inside its tensor
  int kind;   // 0 q, 1 k, 2 v, -1 none
};

__device__ __forceinline__ Unit decode(const PrologueParams& p, long long u) {
  Unit x;
  if (u >= p.units) { x.t = 0; x.h = 0; x.kind = -1; return x; }
  x.t = u / p.upt;
  int r = (int)(u - x.t * p.upt);
  if (r < p.Hq) { x.kind = 0; x.h = r; }
  else if (r < p.Hq + p.Hkv) { x.kind = 1; x.h = r - p.Hq; }
  else { x.kind = 2; x.h = r - p.Hq - p.Hkv; }
  return x;
}

__global__ void __launch_bounds__(kThreads) prologue_kernel(const PrologueParams p) {
  const int lane16 = threadIdx.x & 15;
  const long long hw = (long long)blockIdx.x * kHalfWarpsPerCta + (threadIdx.x >> 4);
  const long long stride = (long long)gridDim.x * kHalfWarpsPerCta;
  const int e0 = 4 * lane16;

  for (long long base = hw; base < p.units; base += stride * kUnroll) {
    Unit un[kUnroll];
    Row x[kUnroll];
    uint2 cs[kUnroll][2];
    long long slot[kUnroll];
#pragma unroll
    for (int i = 0; i < kUnroll; ++i) {
      un[i] = decode(p, base + i * stride);
      slot[i] = -1;
      if (un[i].kind < 0) continue;
      const __nv_bfloat16* src = un[i].kind == 0   ? p.q + un[i].t * p.q_st
                                 : un[i].kind == 1 ? p.k + un[i].t * p.k_st
                                                   : p.v + un[i].t * p.v_st;
      src += (long long)un[i].h * kD + e0;
      x[i].a = ld8_stream(src);
      x[i].b = ld8_stream(src + 64);
      if (un[i].kind != 2 && p.norm_rope) {
        long long pos = p.pos[un[i].t];
        const __nv_bfloat16* c = p.cache + pos * kD + e0;
        cs[i][0] = ld8(c);        // cos[4l .. 4l+3]
        cs[i][1] = ld8(c + 64);   // sin[4l .. 4l+3]
      }
      if (un[i].kind != 0 && p.slots) slot[i] = p.slots[un[i].t];
    }
#pragma unroll
    for (int i = 0; i < kUnroll; ++i) {
      if (un[i].kind < 0) continue;   // uniform per half-warp; shuffles below use the half-warp mask
      const unsigned hmask = 0xffffu << (threadIdx.x & 16);
      if (un[i].kind != 2 && p.norm_rope) {
        float x1[4], x2[4];
        unpack4(x[i].a, x1);
        unpack4(x[i].b, x2);
        float ss = 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) ss += x1[e] * x1[e];   // squares of bf16 are exact in fp32
#pragma unroll
        for (int e = 0; e < 4; ++e) ss += x2[e] * x2[e];
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) ss += __shfl_xor_sync(hmask, ss, o);
        // torch.rsqrt on the CPU is 1 / sqrt(x), both correctly rounded; so are these
        const float rs = __frcp_rn(__fsqrt_rn(ss * (1.0f / kD) + p.eps));
        const __nv_bfloat16* w = (un[i].kind == 0 ? p.q_w : p.k_w) + e0;
        float w1[4], w2[4], c[4], s[4];
        unpack4(ld8(w), w1);
        unpack4(ld8(w + 64), w2);
        unpack4(cs[i][0], c);
        unpack4(cs[i][1], s);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float n1 = rbf(rbf(x1[e] * rs) * w1[e]);   // (x * rsqrt).to(bf16) * weight
          const float n2 = rbf(rbf(x2[e] * rs) * w2[e]);
          x1[e] = rbf(rbf(n1 * c[e]) - rbf(n2 * s[e]));    // o1 = x1*cos - x2*sin, each op rounded
          x2[e] = rbf(rbf(n2 * c[e]) + rbf(n1 * s[e]));    // o2 = x2*cos + x1*sin
        }
        x[i].a = pack4(x1);
        x[i].b = pack4(x2);
        __nv_bfloat16* dst = un[i].kind == 0 ? p.q_out + un[i].t * p.qo_st : p.k_out + un[i].t * p.ko_st;
        dst += (long long)un[i].h * kD + e0;
        st8(dst, x[i].a);
        st8(dst + 64, x[i].b);
      }
      if (slot[i] >= 0) {   // padding tokens carry slot -1 (cache_kernels.cu convention)
        const long long blk = slot[i] / p.block_size;
        const long long off = slot[i] - blk * p.block_size;
        __nv_bfloat16* dst = (un[i].kind == 1 ? p.k_cache : p.v_cache) + blk * p.blk_st + off * p.tok_st +
                             (long long)un[i].h * p.head_st + e0;
        st8(dst, x[i].a);
        st8(dst + 64, x[i].b);
      }
    }
  }
}

bool aligned8(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 7) == 0; }

int launch(const PrologueParams& p, cudaStream_t s) {
  reset_launch_count();
  if (p.units == 0) return TMX_OK;
  int st = check_arch();
  if (st != TMX_OK) return st;
  const long long per_cta = (long long)kHalfWarpsPerCta * kUnroll;
  long long want = (p.units + per_cta - 1) / per_cta;
  const long long cap = (long long)device_sm_count() * 8 * 4;   // 8 resident CTAs/SM x 4 waves
  unsigned grid = (unsigned)(want < cap ? want : cap);
  prologue_kernel<<<grid, kThreads, 0, s>>>(p);
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
    if (sdt % 4 != 0) return TMX_ERR_ALIGNMENT;
  const void* ptrs[] = {q_out, k_out, q, k, q_weight, k_weight, cos_sin_cache, v, k_cache, v_cache};
  for (const void* ptr : ptrs)
    if (!aligned8(ptr)) return TMX_ERR_ALIGNMENT;
  PrologueParams p{};
  p.q = (const __nv_bfloat16*)q; p.k = (const __nv_bfloat16*)k; p.v = (const __nv_bfloat16*)v;
  p.q_out = (__nv_bfloat16*)q_out; p.k_out = (__nv_bfloat16*)k_out;
  p.q_w = (const __nv_bfloat16*)q_weight; p.k_w = (const __nv_bfloat16*)k_weight;
  p.pos = (const long long*)positions; p.cache = (const __nv_bfloat16*)cos_sin_cache;
  p.k_cache = (__nv_bfloat16*)k_cache; p.v_cache = (__nv_bfloat16*)v_cache;
  p.slots = cache ? (const long long*)slot_mapping : nullptr;
  p.q_st = q_stride_t; p.k_st = k_stride_t; p.v_st = v_stride_t; p.qo_st = qo_stride_t; p.ko_st = ko_stride_t;
  p.blk_st = cache_block_stride; p.tok_st = cache_token_stride; p.head_st = cache_head_stride;
  p.T = T; p.Hq = Hq; p.Hkv = Hkv; p.upt = Hq + Hkv + (cache ? Hkv : 0);
  p.units = (long long)T * p.upt;
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
    if (sdt % 4 != 0) return TMX_ERR_ALIGNMENT;
  if (!aligned8(key) || !aligned8(value) || !aligned8(k_cache) || !aligned8(v_cache)) return TMX_ERR_ALIGNMENT;
  PrologueParams p{};
  p.k = (const __nv_bfloat16*)key; p.v = (const __nv_bfloat16*)value;
  p.k_cache = (__nv_bfloat16*)k_cache; p.v_cache = (__nv_bfloat16*)v_cache;
  p.slots = (const long long*)slot_mapping;
  p.k_st = k_stride_t; p.v_st = v_stride_t;
  p.blk_st = cache_block_stride; p.tok_st = cache_token_stride; p.head_st = cache_head_stride;
  p.T = T; p.Hq = 0; p.Hkv = Hkv; p.upt = 2 * Hkv;
  p.units = (long long)T * p.upt;
  p.block_size = block_size; p.norm_rope = 0;
  return launch(p, (cudaStream_t)stream);
}

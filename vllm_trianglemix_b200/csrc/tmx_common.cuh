// Shared device helpers of libtmx: mask predicates in closed form, tile
// classification, error plumbing.  Predicates follow the reference's loops
// (vllm/attention/trianglemix_attention.py:82-89,126-129,166-169,192-196,218).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include "tmx.h"

namespace tmx {

struct MaskParams {
  int si, sl, last, mode;  // mode: TMX_MODE_*
};

// First query row that attends every earlier key: N - i < last  <=>  i >= N - last + 1.
__host__ __device__ __forceinline__ int first_dense_row(int n, int last) {
  long long r = (long long)n - (long long)last + 1;
  return r < 0 ? 0 : (r > n ? n : (int)r);
}

// Triangle / dense predicate for one element (i, j) of a length-n sequence.
__host__ __device__ __forceinline__ bool attends(int i, int j, int n, int lrow,
                                                 int si, int sl, int mode) {
  if (mode == TMX_MODE_NONE) return j < n;
  if (j > i) return false;
  if (mode == TMX_MODE_DENSE_CAUSAL) return true;
  return (j <= si) | (i - j <= sl) | (i >= lrow);
}

// One of the five pieces.
__host__ __device__ __forceinline__ bool piece_at(int piece, int i, int j, int n,
                                                  int lrow, int si, int sl) {
  if (j > i) return false;
  bool stream = (j <= si) | (i - j <= sl);
  bool tail = i >= lrow;
  switch (piece) {
    case TMX_PIECE_STREAMING: return stream;
    case TMX_PIECE_LAST: return !stream & tail;
    case TMX_PIECE_MIDDLE: return !stream & !tail;
    case TMX_PIECE_TRIANGLE: return stream | tail;
    default: return true;
  }
}

// Bits [lo, hi] (inclusive, clipped to 0..31) of a 32-bit word.
__host__ __device__ __forceinline__ uint32_t bit_span(long long lo, long long hi) {
  if (lo < 0) lo = 0;
  if (hi > 31) hi = 31;
  if (hi < lo) return 0u;
  uint32_t upto_hi = (hi == 31) ? 0xffffffffu : ((1u << (hi + 1)) - 1u);
  return upto_hi & ~((1u << lo) - 1u);
}

// Same in 32-bit arithmetic (|lo|, |hi| < 2^31): the form the attention kernels use per tile.
__host__ __device__ __forceinline__ uint32_t bit_span32(int lo, int hi) {
  lo = lo < 0 ? 0 : lo;
  hi = hi > 31 ? 31 : hi;
  if (hi < lo) return 0u;                                 // from here 0 <= lo <= hi <= 31
  return (0xffffffffu >> (31 - hi)) & (0xffffffffu << lo);
}

// Attention mask of query row i for the 32 key columns [c0, c0 + 32) as a word (bit e = column
// c0 + e), same predicate as attends().  One word per 32 scores replaces 32 predicate
// evaluations in the kernels' partial tiles.
__host__ __device__ __forceinline__ uint32_t mask_word32(int i, int c0, int n, int lrow, int si, int sl,
                                                         int mode) {
  if (mode == TMX_MODE_NONE) return bit_span32(-c0, n - 1 - c0);
  const uint32_t causal = bit_span32(-c0, i - c0);
  if (mode == TMX_MODE_DENSE_CAUSAL || i >= lrow) return causal;
  const int wlo = sl > i ? 0 : i - sl;
  const int shi = si < i ? si : i;
  return causal & (bit_span32(-c0, shi - c0) | bit_span32(wlo - c0, i - c0));
}

#ifdef __CUDACC__
// Keeps the mask arithmetic of PARTIAL tiles inside its branch (`if (!full) { cb = pin_in_branch(...); ... }`):
// without the opaque value the compiler evaluates the four mask words of EVERY tile speculatively -- ~130 of the
// ~700 instructions a softmax warp issues per tile (ncu source page of the pair kernel, round 2) although 99.8 %
// of a dense 128K layer's tiles are full.
__device__ __forceinline__ int pin_in_branch(int v) {
#ifndef TMX_DEV_NO_PIN            // developer A/B switch (make DEV=1 DEVEXTRA=-DTMX_DEV_NO_PIN)
  asm volatile("" : "+r"(v));
#endif
  return v;
}
#endif

// 32 columns [c0, c0+32) of row i of a piece, as a word (LSB = column c0).
// Closed form: each row's attended set is a union of at most two intervals.
__host__ __device__ __forceinline__ uint32_t piece_word(int piece, int i, long long c0,
                                                        int n, int lrow, int si, int sl) {
  long long ii = i;
  uint32_t causal = bit_span(0 - c0, ii - c0);
  if (piece == TMX_PIECE_DENSE) return causal;
  long long sink_hi = (si < ii ? (long long)si : ii);
  uint32_t stream = bit_span(0 - c0, sink_hi - c0) | bit_span(ii - sl - c0, ii - c0);
  stream &= causal;
  bool tail = i >= lrow;
  switch (piece) {
    case TMX_PIECE_STREAMING: return stream;
    case TMX_PIECE_LAST: return tail ? (causal & ~stream) : 0u;
    case TMX_PIECE_MIDDLE: return tail ? 0u : (causal & ~stream);
    default: return tail ? causal : stream;  // TRIANGLE
  }
}

// ---- tile level (block index) ---------------------------------------------
// Touched KV tiles of a query tile with rows [r0, r1] (inclusive, r1 < n):
// at most two half-open tile ranges [lo0,hi0) U [lo1,hi1), disjoint, ascending.
struct TileRanges {
  int lo0, hi0, lo1, hi1;
};

__host__ __device__ __forceinline__ TileRanges touched_tiles(int r0, int r1, int n, int bn,
                                                             int lrow, int si, int sl, int mode) {
  TileRanges t;
  if (mode == TMX_MODE_NONE) {
    t.lo0 = 0; t.hi0 = (n - 1) / bn + 1; t.lo1 = t.hi1 = 0;
    return t;
  }
  int diag = r1 / bn;  // last tile holding a column <= r1
  if (mode == TMX_MODE_DENSE_CAUSAL || r1 >= lrow) {
    t.lo0 = 0; t.hi0 = diag + 1; t.lo1 = t.hi1 = 0;
    return t;
  }
  int sink_hi = (si < r1 ? si : r1) / bn;         // tile of the last sink column
  long long w = (long long)r0 - sl;
  int win_lo = (int)((w < 0 ? 0 : w) / bn);       // tile of the first window column
  if (win_lo <= sink_hi + 1) {                    // ranges touch or overlap: merge
    t.lo0 = 0; t.hi0 = diag + 1; t.lo1 = t.hi1 = 0;
  } else {
    t.lo0 = 0; t.hi0 = sink_hi + 1; t.lo1 = win_lo; t.hi1 = diag + 1;
  }
  return t;
}

// Sufficient (not necessary) test that KV tile columns [c0, c1] are attended by
// every row of [r0, r1]; such tiles skip the per-element predicate.
__host__ __device__ __forceinline__ bool tile_is_full(int r0, int r1, int c0, int c1, int n,
                                                      int lrow, int si, int sl, int mode) {
  if (mode == TMX_MODE_NONE) return c1 < n;
  if (c1 > r0 || c1 >= n) return false;
  if (mode == TMX_MODE_DENSE_CAUSAL) return true;
  return (c1 <= si) || ((long long)r1 - c0 <= sl) || (r0 >= lrow);
}

// ---- math without --use_fast_math (which would let the compiler assume no infinities) ----
__device__ __forceinline__ float fast_exp2(float x) {   // ex2.approx: exp2(-inf) = +0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_rcp(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__host__ __device__ __forceinline__ bool is_neg_inf(float x) {
#ifdef __CUDA_ARCH__
  return __float_as_uint(x) == 0xff800000u;
#else
  return x < 0 && (x - x) != 0.0f && x == x * 2.0f;
#endif
}

// ---- error plumbing ---------------------------------------------------------
void set_cuda_error(cudaError_t e, const char* where);
#define TMX_CUDA_TRY(expr)                                   \
  do {                                                       \
    cudaError_t _e = (expr);                                 \
    if (_e != cudaSuccess) {                                 \
      ::tmx::set_cuda_error(_e, #expr);                      \
      return TMX_ERR_CUDA;                                   \
    }                                                        \
  } while (0)

// Per-device caches (tmx_api.cu), keyed by cudaGetDevice(): one process may drive several GPUs.
int device_sm_count();
int check_arch();  // TMX_OK or TMX_ERR_ARCH / TMX_ERR_CUDA
// Has cudaFuncAttributeMaxDynamicSharedMemorySize >= smem already been set for kernel `slot`
// on the CURRENT device?  (function attributes are per device.)  Records the new size if not.
enum { kAttrV3Dense = 0, kAttrV3Tail, kAttrV3Pair, kAttrMma64, kAttrMma128, kAttrStream3 /* + nW/16 */,
       kAttrGemm = kAttrStream3 + 8, kAttrSlots = kAttrGemm + 4 };
bool smem_attr_cached(int slot, int smem);
#define TMX_SET_SMEM_ATTR(kern, slot, smem)                                                        \
  do {                                                                                             \
    if (!::tmx::smem_attr_cached((slot), (smem)))                                                  \
      TMX_CUDA_TRY(cudaFuncSetAttribute((kern), cudaFuncAttributeMaxDynamicSharedMemorySize, (smem))); \
  } while (0)
void count_launch(int n = 1);
void reset_launch_count();

}  // namespace tmx

// Mask kernels: packed bitmask (closed-form interval words, 16-byte stores),
// dense 0/1 and additive {0,-inf} N x N compat outputs, -inf rewrite, and the
// ballot-based additive -> bitmask conversion.  All are pure HBM-write streams:
// one warp per row, lanes striding over 16-byte chunks, so every warp store
// instruction covers 512 contiguous bytes.
//
// Replaces vllm/attention/trianglemix_attention.py:53-261 and
// vllm/attention/ascend_trianglemix.py:114-141 of the reference.
#include "tmx_common.cuh"

namespace tmx {
namespace {

constexpr int kRowsPerCta = 8;  // 8 warps, one row each

// (32-bit interval words: the 64-bit bit_span() of tmx_common.cuh made this kernel issue-bound,
// ~160 instructions per 16-byte store, 2.76 TB/s at 128K.)
// Every piece of every row is a union of at most two column intervals [a0,a1] U [b0,b1]
// (row-uniform, computed once per warp).  A 128-column chunk that lies inside one interval
// is all ones, one that meets neither is all zeros; only the <= 4 chunks per row holding an
// interval end evaluate words.
__global__ void __launch_bounds__(32 * kRowsPerCta)
mask_bits_kernel(uint4* __restrict__ out, long long row_stride_vec, int n, int lrow, int si,
                 int sl, int piece) {
  long long row = (long long)blockIdx.x * kRowsPerCta + threadIdx.y;
  if (row >= n) return;
  const int i = (int)row;
  uint4* dst = out + row * row_stride_vec;
  const int sink_hi = min(si, i);
  const int win_lo = (int)max(0ll, (long long)i - sl);
  const bool tail = i >= lrow;
  int a0 = 0, a1 = -1, b0 = 0, b1 = -1;                    // empty
  switch (piece) {
    case TMX_PIECE_DENSE: a1 = i; break;
    case TMX_PIECE_STREAMING: a1 = sink_hi; b0 = win_lo; b1 = i; break;
    case TMX_PIECE_TRIANGLE:
      if (tail) { a1 = i; } else { a1 = sink_hi; b0 = win_lo; b1 = i; }
      break;
    case TMX_PIECE_LAST: if (tail) { a0 = sink_hi + 1; a1 = win_lo - 1; } break;
    default: if (!tail) { a0 = sink_hi + 1; a1 = win_lo - 1; } break;   // MIDDLE
  }
  const int nvec = (int)row_stride_vec;
#pragma unroll 4
  for (int vcol = threadIdx.x; vcol < nvec; vcol += 32) {
    const int c0 = vcol * 128, c1 = c0 + 127;
    uint4 w;
    if ((c0 >= a0 && c1 <= a1) || (c0 >= b0 && c1 <= b1)) {
      w = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
    } else if ((c1 < a0 || c0 > a1) && (c1 < b0 || c0 > b1)) {
      w = make_uint4(0u, 0u, 0u, 0u);
    } else {
      w.x = bit_span32(a0 - c0, a1 - c0) | bit_span32(b0 - c0, b1 - c0);
      w.y = bit_span32(a0 - c0 - 32, a1 - c0 - 32) | bit_span32(b0 - c0 - 32, b1 - c0 - 32);
      w.z = bit_span32(a0 - c0 - 64, a1 - c0 - 64) | bit_span32(b0 - c0 - 64, b1 - c0 - 64);
      w.w = bit_span32(a0 - c0 - 96, a1 - c0 - 96) | bit_span32(b0 - c0 - 96, b1 - c0 - 96);
    }
    __stcs(dst + vcol, w);
  }
}

template <typename T> struct Vec16 { static constexpr int kN = 16 / sizeof(T); };

template <typename T> __device__ __forceinline__ T from_float(float x);
template <> __device__ __forceinline__ float from_float<float>(float x) { return x; }
template <> __device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float x) {
  return __float2bfloat16(x);
}
template <> __device__ __forceinline__ __half from_float<__half>(float x) { return __float2half(x); }

// kind 0: 0/1 of `piece`; kind 1: additive (0 where attended, -inf elsewhere)
template <typename T, bool kVec>
__global__ void __launch_bounds__(32 * kRowsPerCta)
mask_dense_kernel(T* __restrict__ out, long long ld, int n, int lrow, int si, int sl, int piece,
                  int kind) {
  long long row = (long long)blockIdx.x * kRowsPerCta + threadIdx.y;
  if (row >= n) return;
  int i = (int)row;
  T* dst = out + row * ld;
  const T on = from_float<T>(kind == 0 ? 1.0f : 0.0f);
  const T off = from_float<T>(kind == 0 ? 0.0f : -INFINITY);
  constexpr int kN = Vec16<T>::kN;
  if (kVec) {
    int nvec = n / kN;
    for (int vc = threadIdx.x; vc < nvec; vc += 32) {
      int c0 = vc * kN;
      // the 32-column word containing c0 (kN divides 32, so a chunk never straddles words)
      uint32_t word = (c0 <= i) ? piece_word(piece, i, (long long)(c0 & ~31), n, lrow, si, sl) : 0u;
      word >>= (c0 & 31);
      alignas(16) T vals[kN];
#pragma unroll
      for (int u = 0; u < kN; ++u) vals[u] = ((word >> u) & 1u) ? on : off;
      __stcs(reinterpret_cast<uint4*>(dst + c0), *reinterpret_cast<const uint4*>(vals));
    }
    for (int c = nvec * kN + threadIdx.x; c < n; c += 32)
      dst[c] = piece_at(piece, i, c, n, lrow, si, sl) ? on : off;
  } else {
    for (int c = threadIdx.x; c < n; c += 32)
      dst[c] = piece_at(piece, i, c, n, lrow, si, sl) ? on : off;
  }
}

template <typename T>
__global__ void rewrite_neg_inf_kernel(T* __restrict__ out, const T* __restrict__ in,
                                       long long numel, float fill) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long stride = (long long)gridDim.x * blockDim.x;
  const T f = from_float<T>(fill);
  for (; idx < numel; idx += stride) {
    T x = in[idx];
    float xf = (float)x;
    out[idx] = is_neg_inf(xf) ? f : x;   // only -inf is rewritten
  }
}

// One warp per (row, group of 32 words): lane l handles word l of the group by
// looping over 32 ballots?  No: lane l loads column (32*w + l); ballot gives word w.
// Lanes then own one word each of a 32-word group and store it coalesced.
template <typename T>
__global__ void __launch_bounds__(32 * kRowsPerCta)
bits_from_additive_kernel(uint32_t* __restrict__ out, long long row_stride_words,
                          const T* __restrict__ in, long long ld, int n, int* bad_count) {
  long long row = (long long)blockIdx.x * kRowsPerCta + threadIdx.y;
  if (row >= n) return;
  const T* src = in + row * ld;
  uint32_t* dst = out + row * row_stride_words;
  int lane = threadIdx.x;
  int bad = 0;
  for (long long wbase = 0; wbase < row_stride_words; wbase += 32) {
    uint32_t mine = 0u;
#pragma unroll 4
    for (int w = 0; w < 32; ++w) {
      long long c = (wbase + w) * 32 + lane;
      bool on = false;
      if (wbase + w < row_stride_words && c < n) {
        float x = (float)src[c];
        on = (x == 0.0f);
        // masked: -inf, or any value <= -1e4 (e.g. the -1e6 that optimize_mask_for_npu writes,
        // reference ascend_trianglemix.py:135-139): exp(score + x - rowmax) is exactly 0 in fp32
        bad += (!on && !(x <= -1.0e4f));
      }
      uint32_t word = __ballot_sync(0xffffffffu, on);
      if (w == lane) mine = word;
    }
    if (wbase + lane < row_stride_words) dst[wbase + lane] = mine;
  }
  if (bad_count) {
    for (int o = 16; o > 0; o >>= 1) bad += __shfl_xor_sync(0xffffffffu, bad, o);
    if (lane == 0 && bad) atomicAdd(bad_count, bad);
  }
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

template <typename T>
int launch_dense(void* out, long long ld, int n, int lrow, int si, int sl, int piece, int kind,
                 cudaStream_t s) {
  dim3 block(32, kRowsPerCta);
  unsigned grid = (unsigned)((n + kRowsPerCta - 1) / kRowsPerCta);
  bool vec = aligned16(out) && ((ld * (long long)sizeof(T)) % 16 == 0);
  if (vec)
    mask_dense_kernel<T, true><<<grid, block, 0, s>>>((T*)out, ld, n, lrow, si, sl, piece, kind);
  else
    mask_dense_kernel<T, false><<<grid, block, 0, s>>>((T*)out, ld, n, lrow, si, sl, piece, kind);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

int dense_dispatch(void* out, int dtype, long long ld, int n, int si, int sl, int last, int piece,
                   int kind, cudaStream_t s) {
  if (n < 0 || ld < n) return TMX_ERR_BAD_ARG;
  if (n == 0) return TMX_OK;
  if (!out) return TMX_ERR_BAD_ARG;
  int lrow = first_dense_row(n, last);
  switch (dtype) {
    case TMX_F32: return launch_dense<float>(out, ld, n, lrow, si, sl, piece, kind, s);
    case TMX_BF16: return launch_dense<__nv_bfloat16>(out, ld, n, lrow, si, sl, piece, kind, s);
    case TMX_F16: return launch_dense<__half>(out, ld, n, lrow, si, sl, piece, kind, s);
    default: return TMX_ERR_BAD_ARG;
  }
}

}  // namespace
}  // namespace tmx

using namespace tmx;

extern "C" int tmx_mask_bits(uint32_t* out, int64_t row_stride_words, int N, int si, int sl,
                             int last, int piece, tmx_stream_t stream) {
  if (N < 0 || piece < 0 || piece > TMX_PIECE_DENSE) return TMX_ERR_BAD_ARG;
  if (N == 0) return TMX_OK;
  if (!out) return TMX_ERR_BAD_ARG;
  if (row_stride_words % 4 != 0 || row_stride_words * 32 < N) return TMX_ERR_BAD_SHAPE;
  if (!aligned16(out)) return TMX_ERR_ALIGNMENT;
  dim3 block(32, kRowsPerCta);
  unsigned grid = (unsigned)((N + kRowsPerCta - 1) / kRowsPerCta);
  mask_bits_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<uint4*>(out), row_stride_words / 4, N, first_dense_row(N, last), si, sl, piece);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

extern "C" int tmx_mask_dense01(void* out, int dtype, int64_t ld, int N, int si, int sl, int last,
                                int piece, tmx_stream_t stream) {
  if (piece < 0 || piece > TMX_PIECE_DENSE) return TMX_ERR_BAD_ARG;
  return dense_dispatch(out, dtype, ld, N, si, sl, last, piece, 0, (cudaStream_t)stream);
}

extern "C" int tmx_mask_additive(void* out, int dtype, int64_t ld, int N, int si, int sl, int last,
                                 int mode, tmx_stream_t stream) {
  if (mode != TMX_MODE_DENSE_CAUSAL && mode != TMX_MODE_TRIANGLE) return TMX_ERR_BAD_ARG;
  int piece = mode == TMX_MODE_TRIANGLE ? TMX_PIECE_TRIANGLE : TMX_PIECE_DENSE;
  return dense_dispatch(out, dtype, ld, N, si, sl, last, piece, 1, (cudaStream_t)stream);
}

extern "C" int tmx_mask_rewrite_neg_inf(void* out, const void* in, int dtype, int64_t numel,
                                        float fill, tmx_stream_t stream) {
  if (numel < 0) return TMX_ERR_BAD_ARG;
  if (numel == 0) return TMX_OK;
  if (!out || !in) return TMX_ERR_BAD_ARG;
  int threads = 256;
  long long want = (numel + threads - 1) / threads;
  unsigned grid = (unsigned)(want < 148 * 16 ? want : 148 * 16);
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case TMX_F32: rewrite_neg_inf_kernel<float><<<grid, threads, 0, s>>>((float*)out, (const float*)in, numel, fill); break;
    case TMX_BF16: rewrite_neg_inf_kernel<__nv_bfloat16><<<grid, threads, 0, s>>>((__nv_bfloat16*)out, (const __nv_bfloat16*)in, numel, fill); break;
    case TMX_F16: rewrite_neg_inf_kernel<__half><<<grid, threads, 0, s>>>((__half*)out, (const __half*)in, numel, fill); break;
    default: return TMX_ERR_BAD_ARG;
  }
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

extern "C" int tmx_mask_bits_from_additive(uint32_t* out, int64_t row_stride_words, const void* in,
                                           int dtype, int64_t ld, int N, int32_t* bad_count,
                                           tmx_stream_t stream) {
  if (N < 0 || ld < N) return TMX_ERR_BAD_ARG;
  if (N == 0) return TMX_OK;
  if (!out || !in) return TMX_ERR_BAD_ARG;
  if (row_stride_words * 32 < N) return TMX_ERR_BAD_SHAPE;
  dim3 block(32, kRowsPerCta);
  unsigned grid = (unsigned)((N + kRowsPerCta - 1) / kRowsPerCta);
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case TMX_F32: bits_from_additive_kernel<float><<<grid, block, 0, s>>>(out, row_stride_words, (const float*)in, ld, N, bad_count); break;
    case TMX_BF16: bits_from_additive_kernel<__nv_bfloat16><<<grid, block, 0, s>>>(out, row_stride_words, (const __nv_bfloat16*)in, ld, N, bad_count); break;
    case TMX_F16: bits_from_additive_kernel<__half><<<grid, block, 0, s>>>(out, row_stride_words, (const __half*)in, ld, N, bad_count); break;
    default: return TMX_ERR_BAD_ARG;
  }
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

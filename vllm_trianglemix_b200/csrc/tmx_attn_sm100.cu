// Masked GQA prefill attention on the 5th-generation tensor cores (tcgen05 / TMEM / TMA).
//
// One CTA = one 128-token query tile x TWO query heads that share a KV head (GQA
// head-group reuse: every K/V tile is fetched once for 256 query rows).  10 warps:
//   warps 0-3  softmax group A (head h0):   one thread per query row, TMEM lanes 0..127
//   warps 4-7  softmax group B (head h0+1)
//   warp  8    TMA producer: Q once, then the K/V tiles the block index lists (rings of stages)
//   warp  9    MMA issuer (one elected thread): S = Q K^T into TMEM, O += P V
// KV tiles are 64 keys.  TMEM (512 columns): S_A[2] | S_B[2] (64 fp32 columns each, double
// buffered) | O_A | O_B (128 columns each).  P (bf16) overwrites the first half of the S
// buffer it came from and is read by the MMA straight from TMEM.  Issue order
//   S_A(0) S_B(0) S_A(1) S_B(1) { PV_A(j) S_A(j+2) PV_B(j) S_B(j+2) }_j
// so a group's next scores are already in TMEM when it finishes a tile: the softmax warps
// never wait for their own MMAs, and the tensor pipe works on one group while the other
// exponentiates.  The running max is only refreshed (and O rescaled in TMEM) when it grows
// by more than 2^8, so the correction is off the common path.  Fully masked KV tiles are
// never loaded; only PARTIAL tiles evaluate the mask predicate.
//
// Replaces vllm/attention/ascend_trianglemix.py:86-106 of the reference for dense-causal
// layers (and for any triangle setting through the same block index).
#include <cuda.h>
#include <cudaTypedefs.h>
#include <mutex>
#include "tmx_attn.cuh"
#include "tmx_sm100_ptx.cuh"

namespace tmx {
namespace {

constexpr int kD = 128;
constexpr int kBM = 128;
constexpr int kBN = 64;
constexpr int kThreads = 320;
constexpr int kQBytes = 2 * 2 * kBM * 128;        // 2 heads x 2 panels of [128 rows][64 d] = 64 KB
constexpr int kKStages = 4, kVStages = 3;
constexpr int kPanelBytes = kBN * 128;            // [64 keys][64 d] bf16 = 8 KB
constexpr int kTileBytes = 2 * kPanelBytes;       // K or V tile = 16 KB
constexpr int kOffK = kQBytes;
constexpr int kOffV = kOffK + kKStages * kTileBytes;
constexpr int kOffBar = kOffV + kVStages * kTileBytes;

// barrier indices
constexpr int bQFull = 0;
constexpr int bKFull = 1;                         // [kKStages]
constexpr int bKEmpty = bKFull + kKStages;        // [kKStages]
constexpr int bVFull = bKEmpty + kKStages;        // [kVStages]
constexpr int bVEmpty = bVFull + kVStages;        // [kVStages]
constexpr int bSFull = bVEmpty + kVStages;        // [2 groups][2 buffers]
constexpr int bPReady = bSFull + 4;               // [2 groups][2 buffers], 4 arrivals (one per warp)
constexpr int bPVDone = bPReady + 4;              // [2]: completes after every PV of the group
constexpr int bOFull = bPVDone + 2;               // [2]: the group's last PV has completed
constexpr int kNumBars = bOFull + 2;
constexpr int kSmem = kOffBar + kNumBars * 8 + 16 + 1024;   // + alignment slack

// TMEM columns
constexpr uint32_t kColS = 0;      // + g * 128 + buf * 64
constexpr uint32_t kColO = 256;    // + g * 128

struct TmaMaps {
  CUtensorMap q, k, v;
};

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

}  // namespace
}  // namespace tmx

#include "tmx_attn_sm100_kernel.cuh"

namespace tmx {
namespace {

PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  });
  return fn;
}

// [rows][cols] bf16 row-major with a row pitch of `pitch_elems`; box = 64 columns x box_rows,
// 128-byte swizzle, out-of-bounds rows read as zero.
int make_map(CUtensorMap* m, const void* base, long long rows, long long cols, long long pitch_elems,
             int box_rows) {
  auto enc = get_encode();
  if (!enc) return TMX_ERR_CUDA;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)pitch_elems * 2};
  cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1u, 1u};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_cuda_error(cudaErrorInvalidValue, "cuTensorMapEncodeTiled");
    return TMX_ERR_CUDA;
  }
  return TMX_OK;
}

}  // namespace

int launch_attn_sm100(const AttnParams& p, int variant, cudaStream_t s) {
  if ((p.Hq / p.Hkv) % 2 != 0) return TMX_ERR_UNSUPPORTED;   // needs two query heads per KV head
  if (!(p.scale_log2 > 0.f)) return TMX_ERR_UNSUPPORTED;     // row max is taken on raw scores
  const int emu = variant & 1, dbg = (variant >> 1) & 3, ping = (variant >> 3) & 1;   // TMX_SM100_VARIANT bits
  void (*kern)(const AttnParams, const TmaMaps) = attn_sm100_kernel<0>;
  if (dbg == 1) kern = attn_sm100_kernel<0, 1>;
  else if (dbg == 3) kern = attn_sm100_kernel<0, 3>;
  else if (dbg == 2) kern = ping ? (emu ? attn_sm100_kernel<2, 2, true> : attn_sm100_kernel<0, 2, true>)
                                 : (emu ? attn_sm100_kernel<2, 2> : attn_sm100_kernel<0, 2>);
  else kern = ping ? (emu ? attn_sm100_kernel<2, 0, true> : attn_sm100_kernel<0, 0, true>)
                   : (emu ? attn_sm100_kernel<2> : attn_sm100_kernel<0>);
  TMX_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  TmaMaps maps;
  int st = make_map(&maps.q, p.q, p.total_T, (long long)p.Hq * kD, p.q_stride, kBM);
  if (st == TMX_OK) st = make_map(&maps.k, p.k, p.total_T, (long long)p.Hkv * kD, p.kv_stride, kBN);
  if (st == TMX_OK) st = make_map(&maps.v, p.v, p.total_T, (long long)p.Hkv * kD, p.kv_stride, kBN);
  if (st != TMX_OK) return st;
  dim3 grid((unsigned)p.n_slots, (unsigned)(p.Hq / 2));
  kern<<<grid, kThreads, kSmem, s>>>(p, maps);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

}  // namespace tmx

// Masked GQA prefill attention on the 5th-generation tensor cores (tcgen05 / TMEM / TMA).
//
// One CTA = one 128-token query tile x TWO query heads that share a KV head (GQA
// head-group reuse: every K/V tile is fetched once for 256 query rows).  10 warps:
//   warps 0-3  softmax group A (head h0):   one thread per query row, TMEM lanes 0..127
//   warps 4-7  softmax group B (head h0+1)
//   warp  8    TMA producer: Q once, then the K/V tiles the block index lists (ring of stages)
//   warp  9    MMA issuer (one elected thread): S = Q K^T into TMEM, O += P V
// TMEM (512 columns): S_A | S_B | O_A | O_B, 128 fp32 columns each.  P (bf16) either
// overwrites the first half of its S columns (kPTmem, A operand read from TMEM) or goes to
// shared memory in the 128B-swizzled K-major layout.  The issue order
//   S_A(0) S_B(0) { PV_A(j) S_A(j+1) PV_B(j) S_B(j+1) }_j
// lets group A's softmax overlap group B's MMAs and vice versa.  The running max is only
// refreshed (and O rescaled in TMEM) when it grows by more than 2^8, so the correction is
// off the common path.  Fully masked KV tiles are never loaded; only PARTIAL tiles evaluate
// the mask predicate.
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
constexpr int kThreads = 320;
constexpr int kQBytes = 2 * 2 * kBM * 128;        // 2 heads x 2 panels of [128 rows][64 d] = 64 KB

template <int BN, bool kPTmem> struct Cfg {
  static constexpr int kStages = (BN == 64) ? 3 : (kPTmem ? 2 : 1);
  static constexpr int kPanelBytes = BN * 128;                 // [BN keys][64 d] bf16
  static constexpr int kTileBytes = 2 * kPanelBytes;           // K or V tile
  static constexpr int kPBytes = kPTmem ? 0 : 2 * (BN / 64) * (kBM * 128);   // 2 heads
  static constexpr int kOffK = kQBytes;
  static constexpr int kOffV = kOffK + kStages * kTileBytes;
  static constexpr int kOffP = kOffV + kStages * kTileBytes;
  static constexpr int kOffBar = kOffP + kPBytes;
  static constexpr int kNumBars = 1 + 4 * kStages + 2 + 2 + 2;
  static constexpr int kSmem = kOffBar + kNumBars * 8 + 16 + 1024;   // + alignment slack
};

// barrier indices
template <int BN, bool kPTmem> struct Bars {
  static constexpr int S = Cfg<BN, kPTmem>::kStages;
  static constexpr int q_full = 0;
  static constexpr int k_full = 1;            // [S]
  static constexpr int v_full = 1 + S;        // [S]
  static constexpr int k_empty = 1 + 2 * S;   // [S]
  static constexpr int v_empty = 1 + 3 * S;   // [S]
  static constexpr int s_full = 1 + 4 * S;    // [2]
  static constexpr int p_ready = 3 + 4 * S;   // [2]
  static constexpr int o_full = 5 + 4 * S;    // [2]
};

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

template <int BN, bool kPTmem>
int launch_t(const AttnParams& p, cudaStream_t s) {
  using C = Cfg<BN, kPTmem>;
  static bool attr_set = false;
  if (!attr_set) {
    TMX_CUDA_TRY(cudaFuncSetAttribute(attn_sm100_kernel<BN, kPTmem>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmem));
    attr_set = true;
  }
  TmaMaps maps;
  int st = make_map(&maps.q, p.q, p.total_T, (long long)p.Hq * kD, p.q_stride, kBM);
  if (st == TMX_OK) st = make_map(&maps.k, p.k, p.total_T, (long long)p.Hkv * kD, p.kv_stride, BN);
  if (st == TMX_OK) st = make_map(&maps.v, p.v, p.total_T, (long long)p.Hkv * kD, p.kv_stride, BN);
  if (st != TMX_OK) return st;
  dim3 grid((unsigned)p.n_slots, (unsigned)(p.Hq / 2));
  attn_sm100_kernel<BN, kPTmem><<<grid, kThreads, C::kSmem, s>>>(p, maps);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

}  // namespace

int sm100_block_n(int variant) { return (variant & 1) ? 64 : 128; }

// variant: bit0 = BN 64 (else 128), bit1 = P through shared memory (else TMEM)
int launch_attn_sm100(const AttnParams& p, int variant, cudaStream_t s) {
  if ((p.Hq / p.Hkv) % 2 != 0) return TMX_ERR_UNSUPPORTED;   // needs two query heads per KV head
  switch (variant & 3) {
    case 0: return launch_t<128, true>(p, s);
    case 1: return launch_t<64, true>(p, s);
    case 2: return launch_t<128, false>(p, s);
    default: return launch_t<64, false>(p, s);
  }
}

}  // namespace tmx

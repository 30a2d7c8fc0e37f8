// The step after attention: o_proj as a row-parallel bf16 GEMM on the 5th-generation
// tensor cores whose epilogue IS the reduce-scatter -- every output tile leaves the SM by TMA store straight
// into the staging buffer of the rank that owns its rows (NVLink / NVSwitch peer memory), tile by tile behind
// the math, instead of a full local result followed by a collective.
//
// Reference: Qwen3Attention.forward's `self.o_proj(attn_output)` (vllm/model_executor/models/qwen3.py:190) =
// RowParallelLinear.forward (vllm/model_executor/layers/linear.py:1197-1359): matmul of this rank's
// [T, Hq/P * D] attention shard with its [hidden, Hq/P * D] weight shard, then tensor_model_parallel_all_reduce
// (:1359).  Here: C_partial = A W^T per rank (this kernel), partial tiles pushed to the owner of their token
// rows (slot = source rank), then tmx_sum_partials adds the P partials in fp32 (reduce-scatter along tokens;
// an all-gather of the [T/P, hidden] results gives the reference's all-reduce value).
//
// Kernel: C[M, N] = A[M, K] . B[N, K]^T, bf16 in, fp32 accumulate in TMEM, bf16 out.  Persistent, one CTA per
// SM, tile 128 x 256, K in steps of 64 through a 4-stage TMA ring (48 KB per stage); 6 warps:
//   warps 0-3  epilogue (TMEM lane quadrants): tcgen05.ld -> bf16 -> swizzled staging panel -> TMA store,
//              64 columns at a time through two 16 KB staging panels
//   warp 4     TMA producer        warp 5     MMA issuer (M128 N256 K16 UTCHMMA, A and B from shared memory)
// Two TMEM accumulators (2 x 256 columns): the epilogue of tile i overlaps the main loop of tile i + 1.
#include "tmx_common.cuh"
#include "tmx_sm100_ptx.cuh"
#include "tmx_tma.cuh"

namespace tmx {
namespace {

constexpr int kBM = 128, kBN = 256, kBK = 64;
constexpr int kThreadsG = 192;
constexpr int kABytes = kBM * kBK * 2;            // 16 KB
constexpr int kBBytes = kBN * kBK * 2;            // 32 KB
constexpr int kStageBytes = kABytes + kBBytes;    // 48 KB
constexpr int kStagesG = 4;
constexpr int kPanelBytes = kBM * 128;            // staging panel: 128 rows x 64 bf16 columns
constexpr int kOffStaging = kStagesG * kStageBytes;
constexpr int kOffBarG = kOffStaging + 2 * kPanelBytes;
constexpr int gFull = 0, gEmpty = kStagesG, gAccFull = 2 * kStagesG, gAccEmpty = gAccFull + 2, kNumBarsG = gAccEmpty + 2;
constexpr int kSmemG = kOffBarG + kNumBarsG * 8 + 16 + 1024;
static_assert(kSmemG <= 232448, "exceeds the shared memory of a CTA");
constexpr int kMaxDest = 8;

struct GemmMaps {
  CUtensorMap a, b;
  CUtensorMap out[kMaxDest];      // one per destination rank (peer-memory pointers); out[0] alone = plain GEMM
};

struct GemmParams {
  int M, N, K;
  int m_tiles, n_tiles;
  int n_dest;                     // destination buffers (1: everything goes to out[0])
  int rows_per_dest;              // token rows owned by one destination (multiple of 128 when n_dest > 1)
  int slot;                       // this rank's slot in a destination's staging buffer [n_slots][rows_per_dest][N]
  int m_rot;                      // row-block rotation of the tile walk: rank r starts with the rows of rank r + 1, so
                                  // that at any moment the P ranks store into P different peers (all ranks walking
                                  // rank 0's rows first funnels every store through ONE GPU's NVLink ingress:
                                  // measured 2.08 ms vs 1.02 ms for matmul + ncclReduceScatter at P = 8)
};

__device__ __forceinline__ uint32_t pack2g(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void tma_store_2d(const void* tmap, uint32_t smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"((uint64_t)tmap), "r"(smem_src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }

__global__ void __launch_bounds__(kThreadsG, 1)
gemm_bf16_kernel(const GemmParams p, const __grid_constant__ GemmMaps maps) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sBar = smem_base + kOffBarG;
  const uint32_t sTmemPtr = sBar + kNumBarsG * 8;
  auto bar = [&](int i) { return sBar + 8u * (uint32_t)i; };
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int k_steps = p.K / kBK;
  const int n_tiles_total = p.m_tiles * p.n_tiles;

  if (warp == 4 && lane == 0) {
    ptx::prefetch_tmap(&maps.a); ptx::prefetch_tmap(&maps.b);
    for (int s = 0; s < kStagesG; ++s) { ptx::mbar_init(bar(gFull + s), 1); ptx::mbar_init(bar(gEmpty + s), 1); }
    for (int b = 0; b < 2; ++b) { ptx::mbar_init(bar(gAccFull + b), 1); ptx::mbar_init(bar(gAccEmpty + b), 4); }
    ptx::fence_barrier_init();
  }
  if (warp == 5) ptx::tmem_alloc(sTmemPtr, 512);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(sTmemPtr));

  if (warp == 4) {
    // =========================== TMA producer ===========================
    if (ptx::elect_one()) {
      int it = 0;
      for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x) {
        const int m0 = ((tile / p.n_tiles + p.m_rot) % p.m_tiles) * kBM, n0 = (tile % p.n_tiles) * kBN;
        for (int ks = 0; ks < k_steps; ++ks, ++it) {
          const int st = it % kStagesG;
          if (it >= kStagesG) ptx::mbar_wait(bar(gEmpty + st), ((it / kStagesG) - 1) & 1);
          const uint32_t sA = smem_base + st * kStageBytes, sB = sA + kABytes;
          ptx::mbar_expect_tx(bar(gFull + st), kStageBytes);
          ptx::tma_load_2d(sA, &maps.a, bar(gFull + st), ks * kBK, m0);
          ptx::tma_load_2d(sB, &maps.b, bar(gFull + st), ks * kBK, n0);
          ptx::tma_load_2d(sB + 128 * 128, &maps.b, bar(gFull + st), ks * kBK, n0 + 128);
        }
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    // =========================== MMA issuer ===========================
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    constexpr uint32_t idesc = ptx::instr_desc_bf16(kBM, kBN, false);
    int it = 0, local = 0;
    for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x, ++local) {
      const int b = local & 1;
      if (local >= 2) ptx::mbar_wait(bar(gAccEmpty + b), ((local >> 1) - 1) & 1);   // epilogue drained this accumulator
      ptx::tc_fence_after();
      for (int ks = 0; ks < k_steps; ++ks, ++it) {
        const int st = it % kStagesG;
        ptx::mbar_wait(bar(gFull + st), (it / kStagesG) & 1);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          const uint32_t sA = smem_base + st * kStageBytes, sB = sA + kABytes;
          const uint64_t dA = ptx::smem_desc(sA, 16, 1024), dB = ptx::smem_desc(sB, 16, 1024);
#pragma unroll
          for (int kk = 0; kk < kBK / 16; ++kk)
            ptx::mma_ss(tb + (uint32_t)(b * kBN), dA + (uint64_t)((kk * 32) >> 4), dB + (uint64_t)((kk * 32) >> 4), idesc,
                        (ks > 0 || kk > 0) ? 1u : 0u);
          ptx::mma_commit(bar(gEmpty + st));                       // stage free once these MMAs have read it
          if (ks == k_steps - 1) ptx::mma_commit(bar(gAccFull + b));
        }
        __syncwarp();
      }
    }
  } else {
    // =========================== epilogue ===========================
    const int row = warp * 32 + lane;                              // TMEM lane = output row of the tile
    const uint32_t tl = tmem_base + ((uint32_t)(warp * 32) << 16);
    const uint32_t sStage = smem_base + kOffStaging;
    int local = 0, slab_count = 0;
    for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x, ++local) {
      const int b = local & 1;
      const int m0 = ((tile / p.n_tiles + p.m_rot) % p.m_tiles) * kBM, n0 = (tile % p.n_tiles) * kBN;
      // destination: the rank that owns these token rows; our partial goes to slot `slot` of its buffer
      const int dest = p.n_dest > 1 ? m0 / p.rows_per_dest : 0;
      const int drow = p.n_dest > 1 ? p.slot * p.rows_per_dest + (m0 - dest * p.rows_per_dest) : m0;
      ptx::mbar_wait(bar(gAccFull + b), (local >> 1) & 1);
      ptx::tc_fence_after();
#pragma unroll 1
      for (int slab = 0; slab < kBN / 64; ++slab, ++slab_count) {
        const uint32_t panel = sStage + (uint32_t)((slab_count & 1) * kPanelBytes);
        // the TMA store that last read this panel (two slabs ago) must be done with it
        if (threadIdx.x == 0) bulk_wait_read1();
        ptx::named_bar_sync(1, 128);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t r[32];
          ptx::tmem_ld32(tl + (uint32_t)(b * kBN + slab * 64 + half * 32), r);
          ptx::tc_wait_ld();
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int ch = half * 4 + c;                           // 16-byte chunk of the 128-byte panel row
            const uint32_t addr = panel + (uint32_t)(row * 128 + ((ch ^ (row & 7)) << 4));
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr),
                         "r"(pack2g(__uint_as_float(r[c * 8 + 0]), __uint_as_float(r[c * 8 + 1]))),
                         "r"(pack2g(__uint_as_float(r[c * 8 + 2]), __uint_as_float(r[c * 8 + 3]))),
                         "r"(pack2g(__uint_as_float(r[c * 8 + 4]), __uint_as_float(r[c * 8 + 5]))),
                         "r"(pack2g(__uint_as_float(r[c * 8 + 6]), __uint_as_float(r[c * 8 + 7]))) : "memory");
          }
        }
        ptx::fence_proxy_async_smem();
        ptx::named_bar_sync(1, 128);
        if (threadIdx.x == 0) {
          tma_store_2d(&maps.out[dest], panel, n0 + slab * 64, drow);
          ptx::bulk_commit();
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar(gAccEmpty + b));
    }
    if (threadIdx.x == 0) ptx::bulk_wait_all0();                   // peer / global writes complete before exit
  }
  __syncthreads();
  if (warp == 5) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// out[r, :] = sum over slots s of partials[s, r, :]  (fp32 accumulate, bf16 out): the reduction half of the
// reduce-scatter.  One thread per 8 columns (16-byte accesses), HBM-bound.
__global__ void __launch_bounds__(256) sum_partials_kernel(__nv_bfloat16* __restrict__ out,
                                                           const __nv_bfloat16* __restrict__ partials, int n_slots,
                                                           long long rows, int n, long long out_stride) {
  const long long vec_per_row = n / 8;
  const long long total = rows * vec_per_row;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long r = idx / vec_per_row, c = (idx % vec_per_row) * 8;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int s = 0; s < n_slots; ++s) {
      const uint4 v = *reinterpret_cast<const uint4*>(partials + ((long long)s * rows + r) * n + c);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(h[e]);
        acc[2 * e] += f.x; acc[2 * e + 1] += f.y;
      }
    }
    uint4 o;
    o.x = pack2g(acc[0], acc[1]); o.y = pack2g(acc[2], acc[3]); o.z = pack2g(acc[4], acc[5]); o.w = pack2g(acc[6], acc[7]);
    *reinterpret_cast<uint4*>(out + r * out_stride + c) = o;
  }
}

}  // namespace
}  // namespace tmx

using namespace tmx;

// C[M, N] = A[M, K] . W[N, K]^T in bf16 (fp32 accumulation).  n_dest == 1: the whole result goes to
// out_ptrs_host[0] ([M, N], row stride out_stride).  n_dest > 1 (fused reduce-scatter): rows
// [d * rows_per_dest, (d + 1) * rows_per_dest) go to out_ptrs_host[d], a [n_slots, rows_per_dest, N] staging
// buffer (row stride N) that may live in a PEER GPU's memory, at slot `slot`.
extern "C" int tmx_linear_bf16(void* const* out_ptrs_host, int n_dest, int slot, int n_slots, int64_t rows_per_dest,
                               const void* a, const void* w, int M, int N, int K, int64_t a_stride,
                               int64_t w_stride, int64_t out_stride, tmx_stream_t stream) {
  if (M < 0 || N <= 0 || K <= 0 || n_dest < 1 || n_dest > kMaxDest || slot < 0 || n_slots < 1 || slot >= n_slots)
    return TMX_ERR_BAD_ARG;
  if (M == 0) return TMX_OK;
  if (!out_ptrs_host || !a || !w) return TMX_ERR_BAD_ARG;
  if (N % kBN != 0 || K % kBK != 0) return TMX_ERR_BAD_SHAPE;          // hidden sizes are multiples of 256 / 64
  if (n_dest > 1 && (rows_per_dest <= 0 || rows_per_dest % kBM != 0 || (int64_t)n_dest * rows_per_dest < M))
    return TMX_ERR_BAD_SHAPE;
  if (a_stride < K || w_stride < K || (a_stride % 8) || (w_stride % 8) || (n_dest == 1 && (out_stride < N || out_stride % 8)))
    return TMX_ERR_ALIGNMENT;
  int st = check_arch();
  if (st != TMX_OK) return st;
  GemmMaps maps;
  st = tma_map_2d(&maps.a, a, M, K, a_stride, kBM);
  if (st == TMX_OK) st = tma_map_2d(&maps.b, w, N, K, w_stride, 128);
  for (int d = 0; d < n_dest && st == TMX_OK; ++d) {
    if (!out_ptrs_host[d]) return TMX_ERR_BAD_ARG;
    if (n_dest == 1) st = tma_map_2d(&maps.out[d], out_ptrs_host[d], M, N, out_stride, kBM);
    else st = tma_map_2d(&maps.out[d], out_ptrs_host[d], (long long)n_slots * rows_per_dest, N, N, kBM);
  }
  if (st != TMX_OK) return st;
  GemmParams p{};
  p.M = M; p.N = N; p.K = K;
  p.m_tiles = (M + kBM - 1) / kBM; p.n_tiles = N / kBN;
  p.n_dest = n_dest; p.rows_per_dest = (int)(n_dest > 1 ? rows_per_dest : M); p.slot = slot;
  p.m_rot = n_dest > 1 ? (int)((((int64_t)(slot + 1) % n_dest) * (rows_per_dest / kBM)) % p.m_tiles) : 0;
  TMX_SET_SMEM_ATTR(gemm_bf16_kernel, kAttrGemm, kSmemG);
  const int tiles = p.m_tiles * p.n_tiles, sms = device_sm_count();
  gemm_bf16_kernel<<<tiles < sms ? tiles : sms, kThreadsG, kSmemG, (cudaStream_t)stream>>>(p, maps);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

extern "C" int tmx_sum_partials(void* out, const void* partials, int n_slots, int64_t rows, int N,
                                int64_t out_stride, tmx_stream_t stream) {
  if (n_slots < 1 || rows < 0 || N <= 0 || N % 8 != 0 || out_stride < N || out_stride % 8) return TMX_ERR_BAD_ARG;
  if (rows == 0) return TMX_OK;
  if (!out || !partials) return TMX_ERR_BAD_ARG;
  const long long total = rows * (N / 8);
  const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  sum_partials_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)out, (const __nv_bfloat16*)partials,
                                                                 n_slots, rows, N, out_stride);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

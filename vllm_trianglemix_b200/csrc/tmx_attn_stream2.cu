// Streaming part of a TriangleMix layer (sink columns + causal sliding window), persistent version.
//
// Same tile as tmx_attn_stream.cu (one M=128 tile = TT = 128/G tokens x all G query heads of a KV
// head, sink tile + window tile, every byte of Q/K/V/O moved once), but that kernel is one CTA
// per tile: a serial chain load -> S -> softmax -> PV -> store of ~7 us, two per SM (TMEM), whose
// fixed cost (launch, TMEM alloc, two commit round trips, 16-byte row stores), not DRAM latency,
// kept it at 52 % of the HBM roofline.  Here ONE CTA per SM walks its tiles through a pipeline:
//   warp 8      TMA producer: ring of kStagesMax stages (Q | K window | V window | K sink | V sink)
//   warp 9      MMA issuer: S(t+1) is issued before PV(t), two TMEM sets (S/P columns + O columns)
//   warps 0-3   softmax + epilogue of even tiles (TMEM set 0)
//   warps 4-7   softmax + epilogue of odd tiles  (TMEM set 1)
// The epilogue converts O to bf16 into the stage's (dead) Q buffer in the TMA layout and writes
// it back with ONE 3-D TMA store per 64-d panel; tiles that straddle the end of the streamed
// rows (dense tail / sequence end) fall back to per-row 16-byte stores of the valid rows.
//
// Replaces, for triangle layers, vllm/attention/ascend_trianglemix.py:86-106 of the reference.
#include "tmx_attn.cuh"
#include "tmx_sm100_ptx.cuh"
#include "tmx_tma.cuh"

namespace tmx {
namespace {

constexpr int kD = 128;
constexpr int kSinkKeys = 16;
constexpr int kMaxWin = 112;
constexpr int kThreads2 = 320;
constexpr int kQBytes = 2 * 128 * 128;           // [128 packed rows][128 d] bf16, two 64-d panels
constexpr int kSinkBytes = kSinkKeys * 256;      // K or V sink tile, two panels
constexpr int kStagesMax = 3;
constexpr int kSmemBudget = 225 * 1024;
constexpr uint32_t kColSink = 0, kColWin = 16, kColOut = 128, kColsPerSet = 256;

// barriers: full[3] empty[3] sfull[2] pready[2] ofull[2] tfree[2]
constexpr int bFull = 0, bEmpty = 3, bSFull = 6, bPReady = 8, bOFull = 10, bTFree = 12, kNumBars2 = 14;

struct Stream2Maps {
  CUtensorMap q, kw, vw, ks, vs, o;
};

__device__ __forceinline__ uint32_t pack2s(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

struct Tile {
  int seq_start, r0, n, stream_end, kvh;
  bool valid;
};

// Work item L (KV head fastest) -> tile; identical in every role.
__device__ __forceinline__ Tile tile_at(const AttnParams& p, unsigned L, int TT) {
  Tile t;
  t.kvh = (int)(L % (unsigned)p.Hkv);
  const SlotInfo s = p.slots[L / (unsigned)p.Hkv];
  t.valid = s.seq >= 0;
  t.n = s.n; t.r0 = s.q_tile * TT;
  t.seq_start = t.valid ? p.cu[s.seq] : 0;
  const int lrow = first_dense_row(t.n, p.last);
  t.stream_end = lrow < t.n ? (lrow / p.tail_bm) * p.tail_bm : t.n;
  t.valid = t.valid && t.r0 < t.stream_end;
  return t;
}

__global__ void __launch_bounds__(kThreads2, 1)
attn_stream2_kernel(const AttnParams p, const __grid_constant__ Stream2Maps maps, const int G, const int nW,
                    const int n_stages, const unsigned n_items) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const int winBytes = nW * 256;                              // K or V window tile (two panels)
  const int winPanel = nW * 128;
  const uint32_t stageBytes = (uint32_t)(kQBytes + 2 * winBytes + 2 * kSinkBytes);
  const uint32_t sBar = smem_base + (uint32_t)n_stages * stageBytes;
  const uint32_t sTmemPtr = sBar + kNumBars2 * 8;
  auto bar = [&](int i) { return sBar + 8u * (uint32_t)i; };
  auto stage_base = [&](int st) { return smem_base + (uint32_t)st * stageBytes; };
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int TT = 128 / G;

  if (warp == 8 && lane == 0) {
    ptx::prefetch_tmap(&maps.q); ptx::prefetch_tmap(&maps.kw); ptx::prefetch_tmap(&maps.vw);
    ptx::prefetch_tmap(&maps.ks); ptx::prefetch_tmap(&maps.vs); ptx::prefetch_tmap(&maps.o);
    for (int s = 0; s < kStagesMax; ++s) { ptx::mbar_init(bar(bFull + s), 1); ptx::mbar_init(bar(bEmpty + s), 1); }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(bar(bSFull + s), 1); ptx::mbar_init(bar(bPReady + s), 4);
      ptx::mbar_init(bar(bOFull + s), 1); ptx::mbar_init(bar(bTFree + s), 4);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 9) ptx::tmem_alloc(sTmemPtr, 512);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(sTmemPtr));

  if (warp == 8) {
    // =========================== TMA producer ===========================
    if (ptx::elect_one()) {
      int t = 0;
      for (unsigned L = blockIdx.x; L < n_items; L += gridDim.x) {
        const Tile tl = tile_at(p, L, TT);
        if (!tl.valid) continue;
        const int st = t % n_stages;
        if (t >= n_stages) ptx::mbar_wait(bar(bEmpty + st), ((t / n_stages) - 1) & 1);
        const uint32_t sQ = stage_base(st), sKw = sQ + kQBytes, sVw = sKw + winBytes;
        const uint32_t sKs = sVw + winBytes, sVs = sKs + kSinkBytes;
        const uint32_t full = bar(bFull + st);
        const int w0 = tl.r0 - p.sl, tok = tl.seq_start + tl.r0;
        ptx::mbar_expect_tx(full, stageBytes);
        for (int pnl = 0; pnl < 2; ++pnl) {
          ptx::tma_load_3d(sQ + pnl * (128 * 128), &maps.q, full, pnl * 64, tl.kvh * G, tok);
          ptx::tma_load_2d(sKs + pnl * (kSinkKeys * 128), &maps.ks, full, tl.kvh * kD + pnl * 64, tl.seq_start);
          ptx::tma_load_2d(sKw + pnl * winPanel, &maps.kw, full, tl.kvh * kD + pnl * 64, tl.seq_start + w0);
          ptx::tma_load_2d(sVs + pnl * (kSinkKeys * 128), &maps.vs, full, tl.kvh * kD + pnl * 64, tl.seq_start);
          ptx::tma_load_2d(sVw + pnl * winPanel, &maps.vw, full, tl.kvh * kD + pnl * 64, tl.seq_start + w0);
        }
        ++t;
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // =========================== MMA issuer ===========================
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t idS16 = ptx::instr_desc_bf16(128, kSinkKeys, false);
    const uint32_t idSW = ptx::instr_desc_bf16(128, nW, false);
    const uint32_t idO = ptx::instr_desc_bf16(128, 128, true);
    auto issue_s = [&](int t) {           // scores of local tile t into TMEM set t & 1
      const int st = t % n_stages, set = t & 1;
      ptx::mbar_wait(bar(bFull + st), (t / n_stages) & 1);
      if (t >= 2) ptx::mbar_wait(bar(bTFree + set), ((t >> 1) - 1) & 1);
      ptx::tc_fence_after();
      if (ptx::elect_one()) {
        const uint32_t sQ = stage_base(st), sKw = sQ + kQBytes, sKs = sKw + 2 * winBytes;
        const uint64_t dQ = ptx::smem_desc(sQ, 16, 1024);
        const uint64_t dKs = ptx::smem_desc(sKs, 16, 1024), dKw = ptx::smem_desc(sKw, 16, 1024);
        const uint32_t tc = tb + (uint32_t)set * kColsPerSet;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t qo = ((ks >> 2) * (128 * 128) + (ks & 3) * 32) >> 4;
          ptx::mma_ss(tc + kColSink, dQ + qo, dKs + (((ks >> 2) * (kSinkKeys * 128) + (ks & 3) * 32) >> 4), idS16, ks > 0);
          ptx::mma_ss(tc + kColWin, dQ + qo, dKw + (((ks >> 2) * winPanel + (ks & 3) * 32) >> 4), idSW, ks > 0);
        }
        ptx::mma_commit(bar(bSFull + set));
      }
      __syncwarp();
    };
    // how many valid tiles does this CTA own?  (same walk as the producer)
    int ntiles = 0;
    for (unsigned L = blockIdx.x; L < n_items; L += gridDim.x) ntiles += tile_at(p, L, TT).valid ? 1 : 0;
    ntiles = __shfl_sync(0xffffffffu, ntiles, 0);
    if (ntiles > 0) issue_s(0);
    for (int t = 0; t < ntiles; ++t) {
      if (t + 1 < ntiles) issue_s(t + 1);
      const int st = t % n_stages, set = t & 1;
      ptx::mbar_wait(bar(bPReady + set), (t >> 1) & 1);
      ptx::tc_fence_after();
      if (ptx::elect_one()) {
        const uint32_t sVw = stage_base(st) + kQBytes + winBytes, sVs = sVw + winBytes + kSinkBytes;
        const uint64_t dVs = ptx::smem_desc(sVs, kSinkKeys * 128, 1024), dVw = ptx::smem_desc(sVw, winPanel, 1024);
        const uint32_t tc = tb + (uint32_t)set * kColsPerSet;
        ptx::mma_ts(tc + kColOut, tc, dVs, idO, 0);          // sink keys: P columns [0,8)
        for (int ks = 0; ks < nW / 16; ++ks)
          ptx::mma_ts(tc + kColOut, tc + 8 + (uint32_t)(ks * 8), dVw + (uint64_t)((ks * 2048) >> 4), idO, 1);
        ptx::mma_commit(bar(bOFull + set));
      }
      __syncwarp();
    }
  } else {
#include "tmx_attn_stream2_softmax.inc"
  }

  __syncthreads();
  if (warp == 9) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

// p.slots: one slot per TT-token tile (build_slots with bm = TT), as for launch_attn_stream.
int launch_attn_stream2(const AttnParams& p, cudaStream_t s) {
  const int G = p.Hq / p.Hkv, TT = 128 / G;
  const int nW = (p.sl + TT + 15) / 16 * 16;
  if (nW > kMaxWin) return TMX_ERR_UNSUPPORTED;
  const int stageBytes = kQBytes + 2 * nW * 256 + 2 * kSinkBytes;
  int n_stages = (kSmemBudget - 2048) / stageBytes;
  n_stages = n_stages > kStagesMax ? kStagesMax : n_stages;
  if (n_stages < 2) return TMX_ERR_UNSUPPORTED;
  const int smem = n_stages * stageBytes + kNumBars2 * 8 + 16 + 1024;
  static int smem_set = 0;
  if (smem_set < smem) {
    TMX_CUDA_TRY(cudaFuncSetAttribute(attn_stream2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    smem_set = smem;
  }
  Stream2Maps maps;
  int st = tma_map_q3d(&maps.q, p.q, p.total_T, p.Hq, p.q_stride, G, TT);
  if (st == TMX_OK) st = tma_map_q3d(&maps.o, p.out, p.total_T, p.Hq, p.o_stride, G, TT);
  if (st == TMX_OK) st = tma_map_2d(&maps.kw, p.k, p.total_T, (long long)p.Hkv * kD, p.kv_stride, nW);
  if (st == TMX_OK) st = tma_map_2d(&maps.vw, p.v, p.total_T, (long long)p.Hkv * kD, p.kv_stride, nW);
  if (st == TMX_OK) st = tma_map_2d(&maps.ks, p.k, p.total_T, (long long)p.Hkv * kD, p.kv_stride, kSinkKeys);
  if (st == TMX_OK) st = tma_map_2d(&maps.vs, p.v, p.total_T, (long long)p.Hkv * kD, p.kv_stride, kSinkKeys);
  if (st != TMX_OK) return st;
  const unsigned n_items = (unsigned)p.n_slots * (unsigned)p.Hkv;
  const unsigned sms = (unsigned)device_sm_count();
  dim3 grid(n_items < sms ? n_items : sms);
  attn_stream2_kernel<<<grid, kThreads2, smem, s>>>(p, maps, G, nW, n_stages, n_items);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

}  // namespace tmx

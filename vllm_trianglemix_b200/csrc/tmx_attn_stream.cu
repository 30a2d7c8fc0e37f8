// Streaming part of a TriangleMix layer (sink columns + causal sliding window) on tcgen05.
//
// At the reference's default parameters a triangle layer is HBM-bound (81 FLOP/B): every
// query row attends <= si+1 sink keys and <= sl+1 window keys.  One CTA therefore takes the
// query rows of TT = 128/G consecutive tokens for ALL G query heads of one KV head (GQA
// packing: row m of the M=128 tile = token m/G, head m%G; a 3-D TMA box [TT][G][64] lands
// in exactly the K-major 128-byte-swizzled layout the MMA wants), so that
//   * Q is read once, O written once, and
//   * the K/V rows [r0 - sl, r0 + TT) of the window plus the 16 rows [0, 16) holding the sink
//     are fetched ONCE for the G heads,
// i.e. the CTA moves the algorithmic bytes and nothing else.  Work per CTA is two small
// contractions (S = Q K^T: N = 16 + nW columns, O = P V), one softmax over <= 128 columns per
// row, one 32 KB store -- a short serial chain -- so the kernel is latency-bound per CTA and
// relies on 2-3 resident CTAs per SM (104 KB of shared memory, 256 TMEM columns each).
// Warps 0-3: softmax + epilogue (one thread per packed row), warp 4: TMA, warp 5: MMA issue.
//
// Rows that belong to the dense tail of the sequence (i >= p.stream_end of their sequence:
// the last query tiles, which attend every earlier key) are not stored here; the split-KV
// path writes them.
//
// Replaces, for triangle layers, vllm/attention/ascend_trianglemix.py:86-106 of the reference
// (dense QK^T / softmax / PV over all S x S scores, of which the mask keeps ~38 per row).
#include "tmx_attn.cuh"
#include "tmx_sm100_ptx.cuh"
#include "tmx_tma.cuh"

namespace tmx {
namespace {

constexpr int kD = 128;
constexpr int kSinkKeys = 16;                    // sink tile: keys [0, 16)
constexpr int kMaxWin = 112;                     // window keys per CTA (multiple of 16)
constexpr int kThreadsStream = 192;
constexpr int kQBytes = 2 * 128 * 128;           // [128 packed rows][128 d] bf16, two 64-d panels
constexpr int kOffKw = kQBytes;                  // window K: 2 panels of [nW][64]
constexpr int kOffVw = kOffKw + kMaxWin * 256;
constexpr int kOffKs = kOffVw + kMaxWin * 256;   // sink K: 2 panels of [16][64]
constexpr int kOffVs = kOffKs + kSinkKeys * 256;
constexpr int kOffBarS = kOffVs + kSinkKeys * 256;
constexpr int kSmemStream = kOffBarS + 64 + 1024;
constexpr uint32_t kColSink = 0, kColWin = 16, kColOut = 128;

struct StreamMaps {
  CUtensorMap q, kw, vw, ks, vs;
};

__device__ __forceinline__ uint32_t pack2s(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__global__ void __launch_bounds__(kThreadsStream, 2)
attn_stream_kernel(const AttnParams p, const __grid_constant__ StreamMaps maps, const int G, const int nW,
                   const unsigned pf_stride) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = smem_base, sKw = smem_base + kOffKw, sVw = smem_base + kOffVw;
  const uint32_t sKs = smem_base + kOffKs, sVs = smem_base + kOffVs;
  const uint32_t bIn = smem_base + kOffBarS, bS = bIn + 8, bP = bIn + 16, bO = bIn + 24;
  const uint32_t sTmemPtr = bIn + 32;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  // launch order: the Hkv CTAs of one token tile are neighbours (they read / write adjacent
  // pieces of the same token rows, so DRAM sees whole rows at about the same time)
  const int kvh = (int)(blockIdx.x % (unsigned)p.Hkv);
  const SlotInfo slot = p.slots[blockIdx.x / (unsigned)p.Hkv];
  if (slot.seq < 0) return;
  const int TT = 128 / G;
  const int n = slot.n, r0 = slot.q_tile * TT;
  const int seq_start = p.cu[slot.seq];
  const int w0 = r0 - p.sl;                       // first window key (may be negative: masked)
  // rows >= stream_end belong to query tiles the dense-tail path owns
  const int lrow = first_dense_row(n, p.last);
  const int stream_end = lrow < n ? (lrow / p.tail_bm) * p.tail_bm : n;
  if (r0 >= stream_end) return;

  if (warp == 4 && lane == 0) {
    ptx::prefetch_tmap(&maps.q); ptx::prefetch_tmap(&maps.kw); ptx::prefetch_tmap(&maps.vw);
    ptx::mbar_init(bIn, 1); ptx::mbar_init(bS, 1); ptx::mbar_init(bP, 4); ptx::mbar_init(bO, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 5) ptx::tmem_alloc(sTmemPtr, 256);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(sTmemPtr));
  const int winPanel = nW * 128;                  // bytes of one 64-d panel of the window tiles

  if (warp == 4) {
    if (ptx::elect_one()) {
      ptx::mbar_expect_tx(bIn, (uint32_t)(kQBytes + 2 * 2 * winPanel + 2 * 2 * kSinkKeys * 128));
      for (int pnl = 0; pnl < 2; ++pnl) {
        ptx::tma_load_3d(sQ + pnl * (128 * 128), &maps.q, bIn, pnl * 64, kvh * G, seq_start + r0);
        ptx::tma_load_2d(sKs + pnl * (kSinkKeys * 128), &maps.ks, bIn, kvh * kD + pnl * 64, seq_start);
        ptx::tma_load_2d(sKw + pnl * winPanel, &maps.kw, bIn, kvh * kD + pnl * 64, seq_start + w0);
        ptx::tma_load_2d(sVs + pnl * (kSinkKeys * 128), &maps.vs, bIn, kvh * kD + pnl * 64, seq_start);
        ptx::tma_load_2d(sVw + pnl * winPanel, &maps.vw, bIn, kvh * kD + pnl * 64, seq_start + w0);
      }
      // The CTA is one serial chain (load -> S -> softmax -> PV -> store) and only two fit an SM
      // (TMEM), so HBM latency sits on the critical path.  Pull the inputs of the CTA that will
      // run on this SM slot about one wave from now (same KV head: the stride is a multiple of
      // Hkv) into L2 now; its TMA loads then see L2 latency instead of DRAM latency.
      const unsigned nxt = blockIdx.x + pf_stride;
      if (nxt < gridDim.x) {
        const SlotInfo s2 = p.slots[nxt / (unsigned)p.Hkv];
        if (s2.seq >= 0) {
          const long long t0 = (long long)p.cu[s2.seq] + (long long)s2.q_tile * TT;
          const long long t1 = min(t0 + TT, (long long)p.total_T);
          for (long long tok = t0; tok < t1; ++tok) {
            ptx::prefetch_l2(p.q + tok * p.q_stride + (long long)kvh * G * kD, (uint32_t)(G * kD * 2));
            ptx::prefetch_l2(p.k + tok * p.kv_stride + (long long)kvh * kD, (uint32_t)(kD * 2));
            ptx::prefetch_l2(p.v + tok * p.kv_stride + (long long)kvh * kD, (uint32_t)(kD * 2));
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 5) {
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t idS16 = ptx::instr_desc_bf16(128, kSinkKeys, false);
    const uint32_t idSW = ptx::instr_desc_bf16(128, nW, false);
    const uint32_t idO = ptx::instr_desc_bf16(128, 128, true);
    const uint64_t dQ = ptx::smem_desc(sQ, 16, 1024);
    const uint64_t dKs = ptx::smem_desc(sKs, 16, 1024), dKw = ptx::smem_desc(sKw, 16, 1024);
    const uint64_t dVs = ptx::smem_desc(sVs, kSinkKeys * 128, 1024), dVw = ptx::smem_desc(sVw, winPanel, 1024);
    ptx::mbar_wait(bIn, 0);
    ptx::tc_fence_after();
    if (ptx::elect_one()) {
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {            // S = Q K^T: 8 k-steps of 16 over d
        const uint32_t qo = ((ks >> 2) * (128 * 128) + (ks & 3) * 32) >> 4;
        ptx::mma_ss(tb + kColSink, dQ + qo, dKs + (((ks >> 2) * (kSinkKeys * 128) + (ks & 3) * 32) >> 4), idS16, ks > 0);
        ptx::mma_ss(tb + kColWin, dQ + qo, dKw + (((ks >> 2) * winPanel + (ks & 3) * 32) >> 4), idSW, ks > 0);
      }
      ptx::mma_commit(bS);
    }
    __syncwarp();
    ptx::mbar_wait(bP, 0);
    ptx::tc_fence_after();
    if (ptx::elect_one()) {
      // O = P V: first the 16 sink keys (P columns [0,8)), then nW/16 k-steps of window keys
      ptx::mma_ts(tb + kColOut, tb, dVs, idO, 0);
      for (int ks = 0; ks < nW / 16; ++ks)
        ptx::mma_ts(tb + kColOut, tb + 8 + (uint32_t)(ks * 8), dVw + (uint64_t)((ks * 2048) >> 4), idO, 1);
      ptx::mma_commit(bO);
    }
    __syncwarp();
  } else {
    // =========================== softmax + epilogue ===========================
    const int m = warp * 32 + lane;                // packed row: token m / G, head m % G
    const int i = r0 + m / G;
    const int head = kvh * G + m % G;
    const uint32_t tl = tmem_base + ((uint32_t)(warp * 32) << 16);
    const float sc = p.scale_log2;
    const int nWc = nW / 16;
    // attended columns: sink keys [0, shi] that are not already inside the window, and window
    // columns [clo, chi] (column c of the window tile = key w0 + c)
    const int shi = min(min(p.si, i), i - p.sl - 1);
    const int clo = max(0, i - p.sl) - w0, chi = i - w0;
    auto bits_of = [&](int c) -> uint32_t {        // c = -1: sink chunk, else window chunk c
      return c < 0 ? bit_span(0, shi) : bit_span((long long)clo - c * 16, (long long)chi - c * 16);
    };
    auto col_of = [&](int c) -> uint32_t { return c < 0 ? kColSink : kColWin + (uint32_t)(c * 16); };

    ptx::mbar_wait(bS, 0);
    ptx::tc_fence_after();
    float mx = -INFINITY;
#pragma unroll 1
    for (int c = -1; c < nWc; ++c) {
      const uint32_t bits = bits_of(c);
      uint32_t xr[16];
      ptx::tmem_ld16(tl + col_of(c), xr);
      ptx::tc_wait_ld();
#pragma unroll
      for (int e = 0; e < 16; ++e) mx = ((bits >> e) & 1u) ? fmaxf(mx, __uint_as_float(xr[e])) : mx;
    }
    const float msc = (mx == -INFINITY) ? 0.f : mx * sc;
    float l0 = 0.f, l1 = 0.f;
#pragma unroll 1
    for (int c = -1; c < nWc; ++c) {
      const uint32_t bits = bits_of(c);
      uint32_t xr[16], pk[8];
      ptx::tmem_ld16(tl + col_of(c), xr);
      ptx::tc_wait_ld();
#pragma unroll
      for (int e = 0; e < 16; e += 2) {
        const float a = ((bits >> e) & 1u) ? fast_exp2(fmaf(__uint_as_float(xr[e]), sc, -msc)) : 0.f;
        const float b = ((bits >> (e + 1)) & 1u) ? fast_exp2(fmaf(__uint_as_float(xr[e + 1]), sc, -msc)) : 0.f;
        l0 += a; l1 += b;
        pk[e / 2] = pack2s(a, b);
      }
      ptx::tmem_st8(tl + (c < 0 ? 0u : 8u + (uint32_t)(c * 8)), pk);   // P never overtakes unread S columns
    }
    ptx::tc_wait_st();
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(bP);

    const float l = l0 + l1;
    ptx::mbar_wait(bO, 0);
    ptx::tc_fence_after();
    const float inv = l > 0.f ? fast_rcp(l) : 0.f;
    const bool store = i < stream_end;
    __nv_bfloat16* orow = p.out + (long long)(seq_start + i) * p.o_stride + head * kD;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t r[32];
      ptx::tmem_ld32(tl + kColOut + c * 32, r);
      ptx::tc_wait_ld();
      if (store) {
#pragma unroll
        for (int v4 = 0; v4 < 4; ++v4) {
          uint4 o;
          o.x = pack2s(__uint_as_float(r[v4 * 8 + 0]) * inv, __uint_as_float(r[v4 * 8 + 1]) * inv);
          o.y = pack2s(__uint_as_float(r[v4 * 8 + 2]) * inv, __uint_as_float(r[v4 * 8 + 3]) * inv);
          o.z = pack2s(__uint_as_float(r[v4 * 8 + 4]) * inv, __uint_as_float(r[v4 * 8 + 5]) * inv);
          o.w = pack2s(__uint_as_float(r[v4 * 8 + 6]) * inv, __uint_as_float(r[v4 * 8 + 7]) * inv);
          *reinterpret_cast<uint4*>(orow + c * 32 + v4 * 8) = o;
        }
      }
    }
    if (p.lse != nullptr && store)
      p.lse[(long long)head * p.total_T + seq_start + i] = msc * 0.6931471805599453f + logf(l);
    ptx::tc_fence_before();
  }

  __syncthreads();
  if (warp == 5) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 256);
  }
}

}  // namespace

// Can the streaming kernel serve this triangle setting?  (G query heads per KV head packed
// into M = 128, window of one CTA <= 112 keys, sink inside the first 16 keys.)
bool stream_supported(int Hq, int Hkv, int si, int sl) {
  if (Hkv <= 0 || Hq % Hkv != 0) return false;
  const int G = Hq / Hkv;
  if (G != 2 && G != 4 && G != 8) return false;
  const int nW = (sl + 128 / G + 15) / 16 * 16;
  return si >= 0 && si < kSinkKeys && sl >= 0 && nW <= kMaxWin;
}
int stream_block_m(int Hq, int Hkv) { return 128 / (Hq / Hkv); }

// p.slots: one slot per TT-token tile (build_slots with bm = TT); p.tail_bm: tile size of the
// dense-tail path that owns the rows from its first tail tile on.
int launch_attn_stream(const AttnParams& p, cudaStream_t s) {
  const int G = p.Hq / p.Hkv, TT = 128 / G;
  const int nW = (p.sl + TT + 15) / 16 * 16;
  static bool attr_set = false;
  if (!attr_set) {
    TMX_CUDA_TRY(cudaFuncSetAttribute(attn_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemStream));
    attr_set = true;
  }
  StreamMaps maps;
  int st = tma_map_q3d(&maps.q, p.q, p.total_T, p.Hq, p.q_stride, G, TT);
  if (st == TMX_OK) st = tma_map_2d(&maps.kw, p.k, p.total_T, (long long)p.Hkv * kD, p.kv_stride, nW);
  if (st == TMX_OK) st = tma_map_2d(&maps.vw, p.v, p.total_T, (long long)p.Hkv * kD, p.kv_stride, nW);
  if (st == TMX_OK) st = tma_map_2d(&maps.ks, p.k, p.total_T, (long long)p.Hkv * kD, p.kv_stride, kSinkKeys);
  if (st == TMX_OK) st = tma_map_2d(&maps.vs, p.v, p.total_T, (long long)p.Hkv * kD, p.kv_stride, kSinkKeys);
  if (st != TMX_OK) return st;
  dim3 grid((unsigned)p.n_slots * (unsigned)p.Hkv);
  // CTAs resident at a time (2 per SM), rounded to a multiple of Hkv: the L2 prefetch distance
  unsigned pf = 2u * (unsigned)device_sm_count();
  pf = (pf + (unsigned)p.Hkv - 1) / (unsigned)p.Hkv * (unsigned)p.Hkv;
  attn_stream_kernel<<<grid, kThreadsStream, kSmemStream, s>>>(p, maps, G, nW, pf);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

}  // namespace tmx

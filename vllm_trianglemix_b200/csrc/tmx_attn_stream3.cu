// Streaming part of a TriangleMix layer (sink columns + causal sliding window): the rows before
// the dense tail, which at the reference's default parameters move 2.68 GB per 128K-token layer
// for 0.08 TFLOP -- an HBM-bound pass.
//
// One M = 128 tile = TT = 128/G tokens x all G query heads of a KV head (packed row m = token
// m / G, head m % G, loaded by ONE 3-D TMA box per 64-d panel), against the 16-key sink tile and
// a window tile of nW = 16 * NWC keys: every byte of Q/K/V/O crosses HBM once.  One persistent
// CTA per SM walks its tiles through a pipeline of 12 warps:
//   warp 8      TMA producer: ring of stages (Q | K window | V window | K sink | V sink)
//   warp 9      score issuer: S(t) = Q K^T as soon as tile t has landed; two TMEM sets (S | P | O columns)
//   warp 11     PV issuer: O(t) = P V as soon as P(t) is stored (an in-order single issuer made PV(t)
//               wait behind the LOADS of tile t+1: 41 % of its time, ncu round 2)
//   warps 0-3   softmax + epilogue of even tiles (TMEM set 0)
//   warps 4-7   softmax + epilogue of odd tiles  (TMEM set 1)
//   warp 10     TMA store: O tiles leave through the stage's dead Q buffer, one 3-D box per panel
// Round-2 rewrite of the round-1 kernel (617 us at 128K = 4.3 TB/s).  ncu showed that kernel
// issue-latency-bound (34 % issue slots, long-scoreboard stalls on TMEM loads):
//   * its softmax walked the scores twice in rolled 16-column `tcgen05.ld -> wait -> math`
//     steps (10 exposed round trips per tile); here the whole row (16 + nW <= 128 scores) is
//     loaded once, with one wait, and both passes run from registers;
//   * its MMA warp waited for a "TMEM set free" barrier before S(t+1), which serialised one
//     group's PV behind the OTHER group's epilogue; the barrier is redundant (tcgen05 ops of one
//     thread execute in issue order, and "P(t) ready" already implies that the group finished
//     the epilogue of t-2), so it is gone;
//   * the thread that issued the TMA store blocked its softmax warp in wait_group.read before
//     releasing the stage; a dedicated store warp does that now.
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
constexpr int kThreads3 = 384;
constexpr int kQBytes = 2 * 128 * 128;           // [128 packed rows][128 d] bf16, two 64-d panels
constexpr int kMaxQ = 3, kMaxKV = 4;              // ring depths (Q slots / K-V-O slots)
constexpr int kSmemBudget = 226 * 1024;
constexpr uint32_t kColSink = 0, kColWin = 16, kColOut = 128, kColsPerSet = 256;
// bf16 P: its own columns [80, 120) while the scores fit below them (nW <= 64), so that the next
// scores of a set may be issued as soon as the softmax has READ the old ones; wider windows overlay
// P on the score columns and the next scores wait for PV.
constexpr uint32_t kColPSep = 80;

// barriers: qfull[3] qempty[3] kvfull[4] kvempty[4] staged[4] sfull[2] pready[2] ofull[2]
constexpr int bQFull = 0, bQEmpty = bQFull + kMaxQ, bKVFull = bQEmpty + kMaxQ, bKVEmpty = bKVFull + kMaxKV,
              bStaged = bKVEmpty + kMaxKV, bSFull = bStaged + kMaxKV, bPReady = bSFull + 2, bOFull = bPReady + 2,
              kNumBars3 = bOFull + 2;

struct Stream3Maps {
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

// Developer build: clock64 timeline of the first tiles of a few CTAs (tools/stream_trace.py).
#ifdef TMX_DEV
#define TMX_TRACE(ev, t)                                                                              \
  do {                                                                                                \
    if (p.trace != nullptr && blockIdx.x < 4 && (t) < 96)                                             \
      p.trace[((long long)blockIdx.x * 96 + (t)) * 12 + (ev)] = clock64();                            \
  } while (0)
#else
#define TMX_TRACE(ev, t) do { } while (0)
#endif

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

// Two rings of shared-memory slots: Q (32 KB, free again as soon as the score MMA has read it) and K|V (free
// when the tile's O, staged over the dead K/V bytes, has left by TMA store).  A tile's Q slot lives ~4800 clk, its
// K/V/O slot ~9700 (profiles/r02_stream3_experiments.md), so 2 + 4 slots keep four tiles in flight per SM in the
// 224 KB that three whole-tile stages used to take.
template <int NWC>
__global__ void __launch_bounds__(kThreads3, 1)
attn_stream3_kernel(const AttnParams p, const __grid_constant__ Stream3Maps maps, const int G,
                    const int n_q, const int n_kv, const unsigned n_items) {
  constexpr int nW = NWC * 16;
  // K (and V) of a tile: two 64-d panels of [16 sink keys | nW window keys] rows of 128 bytes, so that the
  // scores of sink + window are ONE N = 16 + nW MMA per k-step and PV walks 1 + NWC uniform 16-key steps
  constexpr int kvPanel = (kSinkKeys + nW) * 128;
  constexpr int kvBytes = 2 * kvPanel;                         // K or V of one tile
  constexpr uint32_t kvSlot = (uint32_t)(2 * kvBytes > kQBytes ? 2 * kvBytes : kQBytes);   // K | V, or the staged O tile
  constexpr bool kSepP = NWC <= 4;
  constexpr uint32_t kColP = kSepP ? kColPSep : 0u;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sKV = smem_base + (uint32_t)n_q * kQBytes;
  const uint32_t sBar = sKV + (uint32_t)n_kv * kvSlot;
  const uint32_t sTmemPtr = sBar + kNumBars3 * 8;
  auto bar = [&](int i) { return sBar + 8u * (uint32_t)i; };
  auto q_slot = [&](int t) { return smem_base + (uint32_t)(t % n_q) * kQBytes; };
  auto kv_slot = [&](int t) { return sKV + (uint32_t)(t % n_kv) * kvSlot; };
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int TT = 128 / G;

  if (warp == 8 && lane == 0) {
    ptx::prefetch_tmap(&maps.q); ptx::prefetch_tmap(&maps.kw); ptx::prefetch_tmap(&maps.vw);
    ptx::prefetch_tmap(&maps.ks); ptx::prefetch_tmap(&maps.vs); ptx::prefetch_tmap(&maps.o);
    for (int s = 0; s < kMaxQ; ++s) { ptx::mbar_init(bar(bQFull + s), 1); ptx::mbar_init(bar(bQEmpty + s), 1); }
    for (int s = 0; s < kMaxKV; ++s) {
      ptx::mbar_init(bar(bKVFull + s), 1); ptx::mbar_init(bar(bKVEmpty + s), 1); ptx::mbar_init(bar(bStaged + s), 4);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(bar(bSFull + s), 1); ptx::mbar_init(bar(bPReady + s), 4); ptx::mbar_init(bar(bOFull + s), 1);
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
        const int qs = t % n_q, ks = t % n_kv;
        if (t >= n_q) ptx::mbar_wait(bar(bQEmpty + qs), ((t / n_q) - 1) & 1);
        if (t >= n_kv) ptx::mbar_wait(bar(bKVEmpty + ks), ((t / n_kv) - 1) & 1);
        TMX_TRACE(0, t);
        const uint32_t sQ = q_slot(t), sK = kv_slot(t), sV = sK + kvBytes;
        const uint32_t qfull = bar(bQFull + qs), kvfull = bar(bKVFull + ks);
        const int w0 = tl.r0 - p.sl, tok = tl.seq_start + tl.r0;
        ptx::mbar_expect_tx(qfull, kQBytes);
        ptx::mbar_expect_tx(kvfull, 2 * kvBytes);
        for (int pnl = 0; pnl < 2; ++pnl) {
          ptx::tma_load_3d(sQ + pnl * (128 * 128), &maps.q, qfull, pnl * 64, tl.kvh * G, tok);
          ptx::tma_load_2d(sK + pnl * kvPanel, &maps.ks, kvfull, tl.kvh * kD + pnl * 64, tl.seq_start);
          ptx::tma_load_2d(sK + pnl * kvPanel + kSinkKeys * 128, &maps.kw, kvfull, tl.kvh * kD + pnl * 64, tl.seq_start + w0);
          ptx::tma_load_2d(sV + pnl * kvPanel, &maps.vs, kvfull, tl.kvh * kD + pnl * 64, tl.seq_start);
          ptx::tma_load_2d(sV + pnl * kvPanel + kSinkKeys * 128, &maps.vw, kvfull, tl.kvh * kD + pnl * 64, tl.seq_start + w0);
        }
        ++t;
      }
    }
    __syncwarp();
  } else if (warp == 9 || warp == 11) {
    // =========================== MMA issuers ===========================
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    constexpr uint32_t idS = ptx::instr_desc_bf16(128, kSinkKeys + nW, false);
    constexpr uint32_t idO = ptx::instr_desc_bf16(128, 128, true);
    // how many valid tiles does this CTA own?  (same walk as the producer)
    int ntiles = 0;
    for (unsigned L = blockIdx.x; L < n_items; L += gridDim.x) ntiles += tile_at(p, L, TT).valid ? 1 : 0;
    ntiles = __shfl_sync(0xffffffffu, ntiles, 0);
    if (warp == 9) {
      // Scores of local tile t into TMEM set t & 1, as soon as (a) the tile has landed and (b) the set's
      // score columns are free: the softmax has read S(t-2) (kSepP: "P(t-2) stored" says so), or, when P
      // overlays the scores, PV(t-2) has consumed it.
      for (int t = 0; t < ntiles; ++t) {
        const int set = t & 1;
        ptx::mbar_wait(bar(bQFull + t % n_q), (t / n_q) & 1);
        ptx::mbar_wait(bar(bKVFull + t % n_kv), (t / n_kv) & 1);
        TMX_TRACE(1, t);
        if (t >= 2) ptx::mbar_wait(bar((kSepP ? bPReady : bOFull) + set), ((t >> 1) - 1) & 1);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          const uint32_t sQ = q_slot(t), sK = kv_slot(t);
          const uint64_t dQ = ptx::smem_desc(sQ, 16, 1024), dK = ptx::smem_desc(sK, 16, 1024);
          const uint32_t tc = tb + (uint32_t)set * kColsPerSet;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            ptx::mma_ss(tc + kColSink, dQ + (((ks >> 2) * (128 * 128) + (ks & 3) * 32) >> 4),
                        dK + (((ks >> 2) * kvPanel + (ks & 3) * 32) >> 4), idS, ks > 0);
          ptx::mma_commit(bar(bSFull + set));
          ptx::mma_commit(bar(bQEmpty + t % n_q));        // Q has been read: its slot may be refilled
          TMX_TRACE(2, t);
        }
        __syncwarp();
      }
    } else {
      for (int t = 0; t < ntiles; ++t) {
        const int set = t & 1;
        // P(t) stored => S(t) was issued (V has landed with it) and this group finished reading O(t-2)
        // (its epilogue precedes its next softmax)
        ptx::mbar_wait(bar(bPReady + set), (t >> 1) & 1);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          const uint64_t dV = ptx::smem_desc(kv_slot(t) + kvBytes, kvPanel, 1024);
          const uint32_t tc = tb + (uint32_t)set * kColsPerSet;
#pragma unroll
          for (int ks = 0; ks <= NWC; ++ks)          // 16 sink keys, then the window in 16-key steps
            ptx::mma_ts(tc + kColOut, tc + kColP + (uint32_t)(ks * 8), dV + (uint64_t)((ks * 2048) >> 4), idO, ks > 0);
          ptx::mma_commit(bar(bOFull + set));
          TMX_TRACE(5, t);
        }
        __syncwarp();
      }
    }
  } else if (warp == 10) {
    // =========================== TMA store ===========================
    // Tiles leave in order; a stage returns to the producer once its O tile has been READ out of
    // shared memory (or at once, for the edge tiles the softmax warps stored row by row).
    if (ptx::elect_one()) {
      int t = 0;
      for (unsigned L = blockIdx.x; L < n_items; L += gridDim.x) {
        const Tile tl = tile_at(p, L, TT);
        if (!tl.valid) continue;
        const int st = t % n_kv;
        ptx::mbar_wait(bar(bStaged + st), (t / n_kv) & 1);
        TMX_TRACE(8, t);
        if (tl.r0 + TT <= tl.stream_end) {
          const uint32_t sO = kv_slot(t);
          ptx::tma_store_3d(&maps.o, sO, 0, tl.kvh * G, tl.seq_start + tl.r0);
          ptx::tma_store_3d(&maps.o, sO + 128 * 128, 64, tl.kvh * G, tl.seq_start + tl.r0);
          ptx::bulk_commit();
          ptx::bulk_wait_read0();
        }
        ptx::mbar_arrive(bar(bKVEmpty + st));
        TMX_TRACE(9, t);
        ++t;
      }
      ptx::bulk_wait_all0();       // global writes complete before the CTA exits
    }
    __syncwarp();
  } else {
    // =========================== softmax + epilogue ===========================
    // group = warp / 4 handles the local tiles t with (t & 1) == group, TMEM set = group.
    const int grp = warp >> 2, w4 = warp & 3;
    const int m = w4 * 32 + lane;                  // packed row: token m / G, head m % G
    const uint32_t tl = tmem_base + ((uint32_t)(w4 * 32) << 16) + (uint32_t)grp * kColsPerSet;
    const float sc = p.scale_log2;
    int t = 0;
    for (unsigned L = blockIdx.x; L < n_items; L += gridDim.x) {
      const Tile tile = tile_at(p, L, TT);
      if (!tile.valid) continue;
      if ((t & 1) != grp) { ++t; continue; }
      const int st = t % n_kv, use = t >> 1;       // `use`: how often this TMEM set was used before
      const int i = tile.r0 + m / G;
      const int head = tile.kvh * G + m % G;
      const int w0 = tile.r0 - p.sl;
      // attended columns: sink keys [0, shi] not already inside the window, window columns [clo, chi]
      const int shi = min(min(p.si, i), i - p.sl - 1);
      const int clo = max(0, i - p.sl) - w0, chi = i - w0;
      const uint32_t sink_bits = bit_span32(0, shi) & 0xffffu;

      ptx::mbar_wait(bar(bSFull + grp), use & 1);
      ptx::tc_fence_after();
      if (m == 0) TMX_TRACE(3, t);
      // the whole row at once: 16 sink scores + nW window scores, ONE wait
      uint32_t xs[16];
      uint32_t xw[NWC][16];
      ptx::tmem_ld16(tl + kColSink, xs);
#pragma unroll
      for (int c = 0; c < NWC; ++c) ptx::tmem_ld16(tl + kColWin + (uint32_t)(c * 16), xw[c]);
      ptx::tc_wait_ld();
      // mask in place (independent selects), then the row max over 8 independent chains
      float mxa[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) mxa[e] = -INFINITY;
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        const float x = ((sink_bits >> e) & 1u) ? __uint_as_float(xs[e]) : -INFINITY;
        xs[e] = __float_as_uint(x);
        mxa[e & 7] = fmaxf(mxa[e & 7], x);
      }
#pragma unroll
      for (int c = 0; c < NWC; ++c) {
        const uint32_t bits = bit_span32(clo - c * 16, chi - c * 16);
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const float x = ((bits >> e) & 1u) ? __uint_as_float(xw[c][e]) : -INFINITY;
          xw[c][e] = __float_as_uint(x);
          mxa[e & 7] = fmaxf(mxa[e & 7], x);
        }
      }
      const float mx = fmaxf(fmaxf(fmaxf(mxa[0], mxa[1]), fmaxf(mxa[2], mxa[3])),
                             fmaxf(fmaxf(mxa[4], mxa[5]), fmaxf(mxa[6], mxa[7])));
      const float msc = (mx == -INFINITY) ? 0.f : mx * sc;     // (the diagonal is always attended)
      float la[4] = {0.f, 0.f, 0.f, 0.f};
      {
        uint32_t pk[8];
#pragma unroll
        for (int e = 0; e < 16; e += 2) {
          const float a = fast_exp2(fmaf(__uint_as_float(xs[e]), sc, -msc));        // exp2(-inf) = 0
          const float b = fast_exp2(fmaf(__uint_as_float(xs[e + 1]), sc, -msc));
          la[(e >> 1) & 3] += a + b;
          pk[e / 2] = pack2s(a, b);
        }
        ptx::tmem_st8(tl + kColP, pk);               // every score is in registers: P may overwrite S
      }
#pragma unroll
      for (int c = 0; c < NWC; ++c) {
        uint32_t pk[8];
#pragma unroll
        for (int e = 0; e < 16; e += 2) {
          const float a = fast_exp2(fmaf(__uint_as_float(xw[c][e]), sc, -msc));
          const float b = fast_exp2(fmaf(__uint_as_float(xw[c][e + 1]), sc, -msc));
          la[(e >> 1) & 3] += a + b;
          pk[e / 2] = pack2s(a, b);
        }
        ptx::tmem_st8(tl + kColP + 8u + (uint32_t)(c * 8), pk);
      }
      ptx::tc_wait_st();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar(bPReady + grp));
      if (m == 0) TMX_TRACE(4, t);

      const float l = (la[0] + la[1]) + (la[2] + la[3]);
      const float inv = l > 0.f ? fast_rcp(l) : 0.f;
      // whole tile inside the streamed rows -> TMA store, staged over the slot's (dead) K | V bytes
      const bool tma_out = tile.r0 + TT <= tile.stream_end;
      const uint32_t sO = kv_slot(t);
      __nv_bfloat16* orow = p.out + (long long)(tile.seq_start + i) * p.o_stride + head * kD;
      if (p.lse != nullptr && i < tile.stream_end)
        p.lse[(long long)head * p.total_T + tile.seq_start + i] = msc * 0.6931471805599453f + logf(l);
      ptx::mbar_wait(bar(bOFull + grp), use & 1);
      ptx::tc_fence_after();
      if (m == 0) TMX_TRACE(6, t);
      auto emit = [&](int c, const uint32_t (&r)[32]) {
#pragma unroll
        for (int v4 = 0; v4 < 4; ++v4) {
          uint4 o;
          o.x = pack2s(__uint_as_float(r[v4 * 8 + 0]) * inv, __uint_as_float(r[v4 * 8 + 1]) * inv);
          o.y = pack2s(__uint_as_float(r[v4 * 8 + 2]) * inv, __uint_as_float(r[v4 * 8 + 3]) * inv);
          o.z = pack2s(__uint_as_float(r[v4 * 8 + 4]) * inv, __uint_as_float(r[v4 * 8 + 5]) * inv);
          o.w = pack2s(__uint_as_float(r[v4 * 8 + 6]) * inv, __uint_as_float(r[v4 * 8 + 7]) * inv);
          if (tma_out) {
            // 128-byte-swizzled K-major panel layout (what the TMA box expects): 16-byte chunk ch of
            // packed row m lives at m * 128 + ((ch ^ (m & 7)) << 4) inside its 64-d panel
            const int d16 = c * 4 + v4;                     // 16-byte chunk of the 256-byte row
            const uint32_t addr = sO + (uint32_t)((d16 >> 3) * (128 * 128) + m * 128 + (((d16 & 7) ^ (m & 7)) << 4));
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
          } else if (i < tile.stream_end) {
            *reinterpret_cast<uint4*>(orow + c * 32 + v4 * 8) = o;
          }
        }
      };
      {
        // O (128 fp32 columns) in four 32-column loads, the next one in flight while one is converted
        uint32_t ra[32], rb[32];
        ptx::tmem_ld32(tl + kColOut, ra);
        ptx::tc_wait_ld();
        ptx::tmem_ld32(tl + kColOut + 32, rb);
        emit(0, ra);
        ptx::tc_wait_ld();
        ptx::tmem_ld32(tl + kColOut + 64, ra);
        emit(1, rb);
        ptx::tc_wait_ld();
        ptx::tmem_ld32(tl + kColOut + 96, rb);
        emit(2, ra);
        ptx::tc_wait_ld();
        emit(3, rb);
      }
      ptx::tc_fence_before();
      if (tma_out) ptx::fence_proxy_async_smem();             // st.shared -> visible to the TMA store
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar(bStaged + st));      // 4 warps: tile staged (or stored row by row)
      if (m == 0) TMX_TRACE(7, t);
      ++t;
    }
  }

  __syncthreads();
  if (warp == 9) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
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

template <int NWC>
static int launch_stream3_t(const AttnParams& p, cudaStream_t s) {
  const int G = p.Hq / p.Hkv, TT = 128 / G;
  constexpr int nW = NWC * 16;
  constexpr int kvBytes2 = 4 * (kSinkKeys + nW) * 128;                     // K | V of one tile
  constexpr int kvSlot = kvBytes2 > kQBytes ? kvBytes2 : kQBytes;          // ... or its staged O tile
  // ring depths: two Q slots, as many K/V/O slots as fit (<= 4)
  const int n_q = 2;
  int n_kv = (kSmemBudget - 2048 - n_q * kQBytes) / kvSlot;
  n_kv = n_kv > kMaxKV ? kMaxKV : n_kv;
  if (n_kv < 2) return TMX_ERR_UNSUPPORTED;
  const int smem = n_q * kQBytes + n_kv * kvSlot + kNumBars3 * 8 + 16 + 1024;
  auto kern = attn_stream3_kernel<NWC>;
  TMX_SET_SMEM_ATTR(kern, kAttrStream3 + NWC, smem);
  Stream3Maps maps;
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
  kern<<<grid, kThreads3, smem, s>>>(p, maps, G, n_q, n_kv, n_items);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

// p.slots: one slot per TT-token tile (build_slots with bm = TT); p.tail_bm: tile size of the
// dense-tail path that owns the rows from its first tail tile on.
int launch_attn_stream3(const AttnParams& p, cudaStream_t s) {
  const int G = p.Hq / p.Hkv, TT = 128 / G;
  const int nwc = (p.sl + TT + 15) / 16;
  switch (nwc) {
    case 1: case 2: return launch_stream3_t<2>(p, s);
    case 3: return launch_stream3_t<3>(p, s);
    case 4: return launch_stream3_t<4>(p, s);
    case 5: return launch_stream3_t<5>(p, s);
    case 6: return launch_stream3_t<6>(p, s);
    case 7: return launch_stream3_t<7>(p, s);
    default: return TMX_ERR_UNSUPPORTED;
  }
}

}  // namespace tmx

// Masked GQA prefill attention on CTA PAIRS (tcgen05 cta_group::2): the dense-layer kernel of round 2.
//
// Round 1's kernel (tmx_attn_sm100_v3.cu) runs one CTA per (128-token query tile, two query heads of a
// KV head).  ncu showed its shared memory saturated: per tile and CTA the tensor pipe reads 192 KB of
// operands (Q and K for the scores, V for PV, twice) and TMA writes 64 KB, against 2048 clk of MMA math.
// Here the two CTAs of a cluster (one TPC) take FOUR query heads of a KV head for the same query tile and
// issue every MMA as one M = 256 instruction across both SMs (UTCHMMA.2CTA):
//   S_g = [Q_g(cta0); Q_g(cta1)] K^T   each CTA holds HALF of the K tile (64 keys): B is split along N
//   O_g += [P_g(cta0); P_g(cta1)] V    each CTA holds HALF of V's columns (one 64-d panel)
// so a CTA loads 32 KB instead of 64 KB per KV tile and the tensor pipe reads 128 KB instead of 192 KB
// of shared memory per tile; K/V leave L2 once per FOUR query heads.
//
// Roles per CTA (10 warps, as in v3): warps 0-3 / 4-7 softmax group A / B (one thread per query row of
// head 2*rank / 2*rank+1), warp 8 TMA producer (its halves; byte counts land on the LEADER's barriers),
// warp 9 MMA issuer (leader CTA only; the peer's warp 9 only takes part in the TMEM allocation).
// TMEM (512 columns per CTA): S_A | S_B | O_A | O_B; bf16 P overwrites columns [0,64) of its S buffer.
// Issue order S_A(0) S_B(0) { PV_A(j) S_A(j+1) PV_B(j) S_B(j+1) }_j; MMA completions are multicast to both
// CTAs' barriers (UTCBAR.2CTA.MULTICAST); "P stored" arrives on the leader's barrier from all 8 warps of a
// group (4 local, 4 remote).
//
// Replaces vllm/attention/ascend_trianglemix.py:86-106 of the reference for dense-causal layers (and any
// block-index triangle setting) when the GQA ratio is a multiple of 4.
#include "tmx_attn.cuh"
#include "tmx_sm100_ptx.cuh"
#include "tmx_tma.cuh"

namespace tmx {
namespace {

constexpr int kD = 128;
constexpr int kBM = 128;
constexpr int kBN = 128;
constexpr int kThreads = 352;                    // 8 softmax warps, producer, two MMA issuers (one per head group)
constexpr int kQBytes = 2 * 2 * kBM * 128;        // 2 heads x 2 panels of [128 rows][64 d] = 64 KB per CTA
constexpr int kKPanel = (kBN / 2) * 128;          // this CTA's 64 keys x 64 d = 8 KB
constexpr int kItemBytes = 2 * kKPanel;           // K half (64 keys x 128 d) or V half (128 keys x 64 d): 16 KB
constexpr int kStages = 10;
constexpr int kOffRing = kQBytes;
constexpr int kOffBar = kOffRing + kStages * kItemBytes;

constexpr int bQFull = 0;
constexpr int bFull = 1;                          // [kStages]  (leader's count the bytes of both CTAs)
constexpr int bEmpty = bFull + kStages;           // [kStages]
constexpr int bSFull = bEmpty + kStages;          // [2 groups]
constexpr int bPReady = bSFull + 2;               // [2 groups], leader: 8 arrivals (4 warps x 2 CTAs)
constexpr int bOFull = bPReady + 2;               // [2 groups]
constexpr int bH1Read = bOFull + 2;               // [2 groups] kAhead: H1's column block has been read (leader: 8 arrivals)
constexpr int bH0Full = bH1Read + 2;              // [2 groups] kAhead: first 64 keys of S(j) complete
constexpr int kNumBars = bH0Full + 2;
constexpr int kSmem = kOffBar + kNumBars * 8 + 16 + 1024;   // + alignment slack
static_assert(kSmem <= 232448, "exceeds the 227 KB of shared memory a CTA may use");

constexpr uint32_t kColS = 0;      // + g * 128
constexpr uint32_t kColO = 256;    // + g * 128

// Developer build: clock64 timeline of the first 96 KV tiles of cluster 1's leader CTA (cluster 0 takes the padding
// slot of a one-sequence batch); tools/pair_trace.py.  Events: 0-3 "H0 seen" of group A's warps 0-3 (one per scheduler),
// 4-7 their "P ready" arrives; issuers: 8-11 H0_A(t+1), PV_A(t) + H1_A(t+1), H0_B(t+1), PV_B(t) + H1_B(t+1) issued.
#ifdef TMX_DEV
#define PAIR_TRACE(ev, t)                                                                                   \
  do {                                                                                                      \
    if (p.trace != nullptr && blockIdx.x == 2 && blockIdx.y == 0 && (t) < 96 && lane == 0)                  \
      p.trace[(long long)(t) * 12 + (ev)] = clock64();                                                      \
  } while (0)
#else
#define PAIR_TRACE(ev, t) do { } while (0)
#endif

struct PairMaps {
  CUtensorMap q, k, v;             // boxes: 128 rows (q), 64 rows (k half), 128 rows (v half)
};

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
// packed fp32 pairs (Blackwell FFMA2 / FADD2)
__device__ __forceinline__ uint64_t pk_f2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk_f2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
// 2^t on the FMA/ALU pipes for a pair (Cody-Waite split + cubic, max rel. error 7.5e-5)
__device__ __forceinline__ uint64_t exp2_poly2(uint64_t t2) {
  float t0, t1;
  upk_f2(t2, t0, t1);
  t2 = pk_f2(fmaxf(t0, -126.0f), fmaxf(t1, -126.0f));
  const uint64_t magic = pk_f2(12582912.0f, 12582912.0f), nmagic = pk_f2(-12582912.0f, -12582912.0f);
  const uint64_t r2 = fadd2(t2, magic);
  const uint64_t u2 = fadd2(r2, nmagic);
  const uint64_t f2 = ffma2(u2, pk_f2(-1.0f, -1.0f), t2);
  uint64_t pl = ffma2(pk_f2(0.0551716685f, 0.0551716685f), f2, pk_f2(0.2426111251f, 0.2426111251f));
  pl = ffma2(pl, f2, pk_f2(0.6932609677f, 0.6932609677f));
  pl = ffma2(pl, f2, pk_f2(0.9999280572f, 0.9999280572f));
  float p0, p1, r0, r1;
  upk_f2(pl, p0, p1);
  upk_f2(r2, r0, r1);
  p0 = __int_as_float(__float_as_int(p0) + (__float_as_int(r0) << 23));
  p1 = __int_as_float(__float_as_int(p1) + (__float_as_int(r1) << 23));
  return pk_f2(p0, p1);
}

// kEmu: how many of every 16 exponentials run on the FMA pipe (even: 0, 4, 6, 8).
// kAhead: the look-ahead schedule of tmx_attn_sm100_v3.cu (kSm == 3) on CTA pairs -- the score MMA of a tile
// is issued as two 64-key halves H0 | H1 into the two 64-column blocks of the group's S buffer, the blocks
// swapping roles every tile (P(j) overwrites H0(j)'s block); H0(j+1) is issued as soon as both CTAs' softmax
// warps have read H1(j) -- before P(j) exists -- and H1(j+1) follows PV(j).  The softmax pass is software-
// pipelined over its four 32-column chunks.  Each CTA then holds keys {32 rank .. +31} of BOTH halves of the
// K tile (four 32-key boxes per tile).  Barriers: bH0Full = "H0 ready", bSFull = "H1 ready", bH1Read.
//
// kSplit (needs kAhead): TWO threads per query row -- 16 softmax warps per CTA instead of 8.  Warp w serves lane
// quadrant w & 3 of group (w >> 2) & 1 and HALF h = w >> 3 of every tile: h = 0 the 64 keys of H0, h = 1 those of
// H1.  The optimistic pass needs no row max, so the two threads of a row only agree on "stale reference?" once per
// tile (one float through shared memory + a 64-thread named barrier); the rare refresh exchanges the half maxima
// the same way and each half rescales its own 64 columns of O.  P(j) still lands in H0's block (h writes columns
// [32h, 32h + 32)), the epilogue splits the 128 output columns.  A warp's pass is half as long and four warps per
// scheduler hide each other's MUFU / TMEM latencies (the 8-warp pass issues at 0.5 IPC per scheduler).
constexpr int kThreadsSplit = 576;
template <int kEmu, bool kAhead, bool kSplit = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kSplit ? kThreadsSplit : kThreads, 1)
attn_sm100_pair_kernel(const AttnParams p, const __grid_constant__ PairMaps maps) {
  static_assert(!kSplit || kAhead, "the split softmax builds on the look-ahead schedule");
  constexpr int kWProd = kSplit ? 16 : 8, kWMma = kSplit ? 17 : 9;     // producer / MMA-issuer warp
  constexpr int kRing = kSplit ? kStages - 1 : kStages;               // kSplit: the last slot is the exchange area
#ifdef TMX_DEV_INORDER_ISSUER                                           // developer A/B switch
  constexpr bool kOoo = false;
#else
  constexpr bool kOoo = kAhead && !kSplit && (kStages % 2 == 0);        // one issuer warp per head group (ring positions advance by 2)
#endif
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;   // SWIZZLE_128B: 1024-byte atoms
  const uint32_t sQ = smem_base, sRing = smem_base + kOffRing;
  const uint32_t sBar = smem_base + kOffBar;
  const uint32_t sTmemPtr = sBar + kNumBars * 8;
  auto bar = [&](int i) { return sBar + 8u * (uint32_t)i; };

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // provably warp-uniform
  const int lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();                           // 0 = leader
  const bool leader = rank == 0;

  // ---- work item: uniform per PAIR (both CTAs read the same slot) ---------------------------
  // r0 / r1 / i are ABSOLUTE rows of the (context + chunk) sequence (see tmx_attn_sm100_v3.cu).
  const SlotInfo slot = p.slots[p.n_slots - 1 - (blockIdx.x >> 1)];        // descending: longest first
  const bool active = slot.seq >= 0;
  int n = 0, r0 = 0, r1 = -1, lo0 = 0, cnt0 = 0, lo1 = 0, total = 0, seq_start = 0, q_row0 = 0, k_start = 0, lrow = 0;
  if (active) {
    const SeqGeom gm = seq_geom(p, slot.seq);
    n = slot.n; r0 = gm.ctx + slot.q_tile * kBM;
    r1 = min(n, r0 + kBM) - 1;
    lrow = first_dense_row(n, p.last);
    if (p.tail_clip && p.mode == TMX_MODE_TRIANGLE && lrow < n) r1 = min(r1, lrow - 1);
    lo0 = slot.lo0; cnt0 = slot.hi0 - slot.lo0; lo1 = slot.lo1;
    total = cnt0 + (slot.hi1 - slot.lo1);
    seq_start = gm.q_start - gm.ctx;      // query-buffer row of absolute row 0 (may be negative)
    q_row0 = gm.q_start + slot.q_tile * kBM; k_start = gm.k_start;
  }
  if (!active) return;                                 // both CTAs of the pair agree
  const int h0 = 4 * blockIdx.y + 2 * (int)rank;      // this CTA's two query heads
  const int kvh = (4 * blockIdx.y) / (p.Hq / p.Hkv);
  auto tile_of = [&](int it) { return it < cnt0 ? lo0 + it : lo1 + (it - cnt0); };

  // ---- setup -------------------------------------------------------------------
  if (warp == kWProd && lane == 0) {
    ptx::prefetch_tmap(&maps.q); ptx::prefetch_tmap(&maps.k); ptx::prefetch_tmap(&maps.v);
    ptx::mbar_init(bar(bQFull), 1);
    for (int s = 0; s < kStages; ++s) { ptx::mbar_init(bar(bFull + s), 1); ptx::mbar_init(bar(bEmpty + s), kOoo ? 2 : 1); }
    for (int g = 0; g < 2; ++g) {
      ptx::mbar_init(bar(bSFull + g), 1);
      ptx::mbar_init(bar(bPReady + g), kSplit ? 16 : 8);   // one arrive per softmax warp of the group, both CTAs
      ptx::mbar_init(bar(bOFull + g), 1);
      ptx::mbar_init(bar(bH1Read + g), 8);
      ptx::mbar_init(bar(bH0Full + g), 1);
    }
    ptx::fence_barrier_init();
  }
  if (warp == kWMma) ptx::tmem_alloc_pair(sTmemPtr, 512);
  ptx::tc_fence_before();
  ptx::cluster_sync();                            // barriers of BOTH CTAs exist before any remote arrive / TMA
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(sTmemPtr));

  if (warp == kWProd) {
    // =========================== TMA producer (both CTAs) ===========================
    // Item t: even = K(t/2), odd = V(t/2).  This CTA's half: keys [64 rank, 64 rank + 64) of the K tile
    // (both d panels), d panel `rank` of the V tile (all 128 keys).  Bytes are counted on the leader's barrier.
    if (ptx::elect_one()) {
      if (leader) ptx::mbar_expect_tx(bar(bQFull), 2 * kQBytes);
      for (int g = 0; g < 2; ++g)
        for (int pnl = 0; pnl < 2; ++pnl)
          ptx::tma_load_2d_pair(sQ + (g * 2 + pnl) * (kBM * 128), &maps.q, bar(bQFull), (h0 + g) * kD + pnl * 64, q_row0);
      const int items = 2 * total;
      int st = 0;
      for (int t = 0; t < items; ++t) {
        if (t >= kRing) ptx::mbar_wait(bar(bEmpty + st), ((t / kRing) - 1) & 1);
        if (leader) ptx::mbar_expect_tx(bar(bFull + st), 2 * kItemBytes);
        const int row = k_start + tile_of(t >> 1) * kBN;
        const uint32_t dst = sRing + st * kItemBytes;
        if ((t & 1) == 0) {
          if constexpr (kAhead) {
            // [half h][panel][32 keys][128 B]: keys 64 h + 32 rank .. + 31 of the tile
            for (int h = 0; h < 2; ++h)
              for (int pnl = 0; pnl < 2; ++pnl)
                ptx::tma_load_2d_pair(dst + h * kKPanel + pnl * (kKPanel / 2), &maps.k, bar(bFull + st),
                                      kvh * kD + pnl * 64, row + 64 * h + 32 * (int)rank);
          } else {
            for (int pnl = 0; pnl < 2; ++pnl)
              ptx::tma_load_2d_pair(dst + pnl * kKPanel, &maps.k, bar(bFull + st), kvh * kD + pnl * 64, row + 64 * (int)rank);
          }
        } else {
          ptx::tma_load_2d_pair(dst, &maps.v, bar(bFull + st), kvh * kD + 64 * (int)rank, row);
        }
        if (++st == kRing) st = 0;
      }
    }
    __syncwarp();
  } else if (warp == kWMma || (kOoo && warp == kWMma + 1)) {
    // =========================== MMA issuer(s) (leader CTA only) ===========================
    if (leader) {
      const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
      const int ntot = __shfl_sync(0xffffffffu, total, 0);
      constexpr uint32_t kIdescS = ptx::instr_desc_bf16(256, kBN, false);
      constexpr uint32_t kIdescO = ptx::instr_desc_bf16(256, 128, true);
      const uint64_t dQ = ptx::smem_desc(sQ, 16, 1024);
      const uint64_t dK = ptx::smem_desc(sRing, 16, 1024);             // K half: K-major (d contiguous)
      const uint64_t dV = ptx::smem_desc(sRing, kItemBytes, 1024);     // V half: MN-major, one 64-d panel

      // S_g = Q_g K^T of the K tile in ring stage `st`: 8 k-steps of 16 over d (two 64-d panels)
      auto issue_s = [&](int g, int st) {
        const uint32_t d_tmem = tb + kColS + (uint32_t)(g * 128);
        const uint64_t qa = dQ + (uint64_t)((g * 2 * (kBM * 128)) >> 4);
        const uint64_t kb = dK + (uint64_t)((st * kItemBytes) >> 4);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t qo = ((ks >> 2) * (kBM * 128) + (ks & 3) * 32) >> 4;
          const uint32_t ko = ((ks >> 2) * kKPanel + (ks & 3) * 32) >> 4;
          ptx::mma_ss_pair(d_tmem, qa + qo, kb + ko, kIdescS, ks > 0);
        }
      };
      // O_g += P_g V: 8 k-steps of 16 keys; P (bf16) read from columns [0,64) of S_g
      auto issue_pv = [&](int g, int st, bool acc) {
        const uint32_t d_tmem = tb + kColO + (uint32_t)(g * 128);
        const uint32_t a_tmem = tb + kColS + (uint32_t)(g * 128);
        const uint64_t vb = dV + (uint64_t)((st * kItemBytes) >> 4);
#pragma unroll
        for (int ks = 0; ks < kBN / 16; ++ks)
          ptx::mma_ts_pair(d_tmem, a_tmem + (uint32_t)(ks * 8), vb + (uint64_t)((ks * 2048) >> 4), kIdescO,
                           acc || ks > 0);
      };

      // kAhead: half h (64 keys) of S_g for tile j into column block (j & 1) ^ h
      constexpr uint32_t kIdescS64 = ptx::instr_desc_bf16(256, 64, false);
      auto issue_s_half = [&](int g, int st, int j, int h) {
        const uint32_t d_tmem = tb + kColS + (uint32_t)(g * 128) + (uint32_t)((((j & 1) ^ h)) * 64);
        const uint64_t qa = dQ + (uint64_t)((g * 2 * (kBM * 128)) >> 4);
        const uint64_t kb = dK + (uint64_t)((st * kItemBytes + h * kKPanel) >> 4);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t qo = ((ks >> 2) * (kBM * 128) + (ks & 3) * 32) >> 4;
          const uint32_t ko = ((ks >> 2) * (kKPanel / 2) + (ks & 3) * 32) >> 4;
          ptx::mma_ss_pair(d_tmem, qa + qo, kb + ko, kIdescS64, ks > 0);
        }
      };
      auto issue_pv_blk = [&](int g, int st, bool acc, int j) {     // P(j) sits in block j & 1
        const uint32_t d_tmem = tb + kColO + (uint32_t)(g * 128);
        const uint32_t a_tmem = tb + kColS + (uint32_t)(g * 128) + (uint32_t)((j & 1) * 64);
        const uint64_t vb = dV + (uint64_t)((st * kItemBytes) >> 4);
#pragma unroll
        for (int ks = 0; ks < kBN / 16; ++ks)
          ptx::mma_ts_pair(d_tmem, a_tmem + (uint32_t)(ks * 8), vb + (uint64_t)((ks * 2048) >> 4), kIdescO,
                           acc || ks > 0);
      };

      int st = 0, ph = 0;                               // ring position of the next item to consume
      auto advance = [&] { if (++st == kRing) { st = 0; ph ^= 1; } };
      ptx::mbar_wait_cluster(bar(bQFull), 0);
      ptx::mbar_wait_cluster(bar(bFull + 0), 0);        // K(0)
      ptx::tc_fence_after();
      if constexpr (kOoo) {
        // One issuer WARP per head group (warp kWMma: group A, kWMma + 1: group B), each walking its own chain
        //   "H1 read"(j) -> H0(j+1) ... "P ready"(j) -> PV(j), H1(j+1)
        // with sleeping mbarrier waits.  History (profiles/r02_pair_kernel.md): a single in-order issuer could not
        // issue H0_A(j+1) while it waited for "P ready" of group B, which locked the two groups' softmax passes to
        // each other; a single POLLING issuer fixed that (+1.2 %) but its poll loop took issue slots from the two
        // softmax warps of its scheduler -- clock64 trace: warp 1 arrived 900 clk after warps 0, 2, 3 at every
        // "P ready", and the M = 256 MMAs wait for the slowest of the 8 warps.  tcgen05 operations of different
        // threads are unordered, which is fine here: the groups share no TMEM columns, and a ring stage is
        // released by one commit per group (count 2), each after that group's last read of it.
        const int g = warp - kWMma;
        if (ptx::elect_one()) {
          issue_s_half(g, 0, 0, 0); ptx::mma_commit_pair(bar(bH0Full + g));
          issue_s_half(g, 0, 0, 1); ptx::mma_commit_pair(bar(bSFull + g));
          ptx::mma_commit_pair(bar(bEmpty + 0));
        }
        __syncwarp();
        int vst = 1, vph = 0, kst = 2, kph = 0;          // ring positions of V(j) and K(j+1): items 2j+1 and 2j+2
        for (int j = 0; j < ntot; ++j) {
          const bool more = j + 1 < ntot;
          if (more) {
            ptx::mbar_wait_cluster(bar(bFull + kst), kph);
            ptx::mbar_wait_cluster(bar(bH1Read + g), j & 1);         // H1(j)'s block has been read (both CTAs)
            ptx::tc_fence_after();
            if (ptx::elect_one()) { issue_s_half(g, kst, j + 1, 0); ptx::mma_commit_pair(bar(bH0Full + g)); }
            __syncwarp();
            PAIR_TRACE(8 + 2 * g, j);
          }
          ptx::mbar_wait_cluster(bar(bFull + vst), vph);
          ptx::mbar_wait_cluster(bar(bPReady + g), j & 1);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            issue_pv_blk(g, vst, j > 0, j);
            if (!more) ptx::mma_commit_pair(bar(bOFull + g));
            ptx::mma_commit_pair(bar(bEmpty + vst));
            if (more) {
              issue_s_half(g, kst, j + 1, 1);                        // into P(j)'s block, behind PV(j)
              ptx::mma_commit_pair(bar(bSFull + g));
              ptx::mma_commit_pair(bar(bEmpty + kst));
            }
          }
          __syncwarp();
          PAIR_TRACE(9 + 2 * g, j);
          vst += 2; if (vst >= kRing) { vst -= kRing; vph ^= 1; }
          kst += 2; if (kst >= kRing) { kst -= kRing; kph ^= 1; }
        }
      } else if constexpr (kAhead) {
        if (ptx::elect_one()) {
          for (int g = 0; g < 2; ++g) {
            issue_s_half(g, 0, 0, 0); ptx::mma_commit_pair(bar(bH0Full + g));
            issue_s_half(g, 0, 0, 1); ptx::mma_commit_pair(bar(bSFull + g));
          }
          ptx::mma_commit_pair(bar(bEmpty + 0));
        }
        __syncwarp();
        advance();
        for (int j = 0; j < ntot; ++j) {
          const int vst = st, vph = ph;
          advance();
          const int kst = st, kph = ph;
          const bool more = j + 1 < ntot;
          if (more) advance();
          ptx::mbar_wait_cluster(bar(bFull + vst), vph);
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            if (more) {
              if (g == 0) ptx::mbar_wait_cluster(bar(bFull + kst), kph);
              ptx::mbar_wait_cluster(bar(bH1Read + g), j & 1);       // H1(j)'s block has been read (both CTAs)
              ptx::tc_fence_after();
              if (ptx::elect_one()) { issue_s_half(g, kst, j + 1, 0); ptx::mma_commit_pair(bar(bH0Full + g)); }
              __syncwarp();
            }
            ptx::mbar_wait_cluster(bar(bPReady + g), j & 1);
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
              issue_pv_blk(g, vst, j > 0, j);
              if (!more) ptx::mma_commit_pair(bar(bOFull + g));
              if (g == 1) ptx::mma_commit_pair(bar(bEmpty + vst));
              if (more) {
                issue_s_half(g, kst, j + 1, 1);                      // into P(j)'s block, behind PV(j)
                ptx::mma_commit_pair(bar(bSFull + g));
                if (g == 1) ptx::mma_commit_pair(bar(bEmpty + kst));
              }
            }
            __syncwarp();
          }
        }
      } else {
      if (ptx::elect_one()) {
        issue_s(0, 0); ptx::mma_commit_pair(bar(bSFull + 0));
        issue_s(1, 0); ptx::mma_commit_pair(bar(bSFull + 1));
        ptx::mma_commit_pair(bar(bEmpty + 0));
      }
      __syncwarp();
      advance();
      for (int j = 0; j < ntot; ++j) {
        const int vst = st, vph = ph;
        advance();
        const int kst = st, kph = ph;
        const bool more = j + 1 < ntot;
        if (more) advance();
        ptx::mbar_wait_cluster(bar(bFull + vst), vph);
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          ptx::mbar_wait_cluster(bar(bPReady + g), j & 1);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            issue_pv(g, vst, j > 0);
            if (!more) ptx::mma_commit_pair(bar(bOFull + g));
            if (g == 1) ptx::mma_commit_pair(bar(bEmpty + vst));
          }
          __syncwarp();
          if (more) {
            if (g == 0) { ptx::mbar_wait_cluster(bar(bFull + kst), kph); ptx::tc_fence_after(); }
            if (ptx::elect_one()) {
              issue_s(g, kst);
              ptx::mma_commit_pair(bar(bSFull + g));
              if (g == 1) ptx::mma_commit_pair(bar(bEmpty + kst));
            }
            __syncwarp();
          }
        }
      }
      }   // !kAhead
    }
  } else if constexpr (kSplit) {
    // =========================== softmax, two threads per row (both CTAs) ===========================
    const int q4 = warp & 3, g = (warp >> 2) & 1, h = warp >> 3;     // lane quadrant, head group, tile half
    const int i = r0 + q4 * 32 + lane;             // query row inside the sequence
    const int head = h0 + g;
    const uint32_t lane_base = (uint32_t)(q4 * 32) << 16;
    const uint32_t tS = tmem_base + lane_base + kColS + (uint32_t)(g * 128);
    const uint32_t tO = tmem_base + lane_base + kColO + (uint32_t)(g * 128) + (uint32_t)(h * 64);   // this half's 64 O columns
    const float sc = p.scale_log2;
    const uint64_t sc2 = pk_f2(sc, sc);
    float m_ref = -INFINITY, l_run = 0.f;          // shared reference max (identical in both halves), own partial sum
    const uint32_t* bm_row = p.bitmask ? p.bitmask + (long long)min(i, n - 1) * p.bitmask_stride : nullptr;
    // exchange with the other half of the same rows: one float through shared memory (the ring's last slot),
    // double-buffered, ordered by a 64-thread named barrier
    const uint32_t sX = sRing + (uint32_t)(kRing * kItemBytes);
    const int bar_id = 1 + g * 4 + q4;
    int xc = 0;
    auto exchange = [&](float mine) -> float {
      const uint32_t base = sX + (uint32_t)((((xc & 1) * 8 + g * 4 + q4) * 2) * 128) + (uint32_t)(lane * 4);
      ++xc;
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(base + (uint32_t)(h * 128)), "f"(mine) : "memory");
      ptx::named_bar_sync(bar_id, 64);
      float other;
      asm volatile("ld.shared.f32 %0, [%1];" : "=f"(other) : "r"(base + (uint32_t)((h ^ 1) * 128)) : "memory");
      return other;
    };

    for (int j = 0; j < total; ++j) {
      const int c0 = tile_of(j) * kBN;
      const bool full = bm_row == nullptr &&
                        tile_is_full(r0, r1, c0, c0 + kBN - 1, n, lrow, p.si, p.sl, p.mode);
      // this half's 64-column block: H0 sits in block j & 1, H1 in the other; P(j) goes to H0's block
      const uint32_t tH0 = tS + (uint32_t)((j & 1) * 64);
      const uint32_t tB = tS + (uint32_t)((((j & 1) ^ h)) * 64);
      auto wait_block = [&]() {
        ptx::mbar_wait(bar((h == 0 ? bH0Full : bSFull) + g), j & 1);
        ptx::tc_fence_after();
      };
      auto masked = [&](int c, uint32_t (&xr)[32]) {          // chunk c (0 / 1) of this half; PARTIAL tiles only
        if (!full) {
          const int cb = pin_in_branch(c0 + h * 64 + c * 32);
          const uint32_t bits = bm_row != nullptr ? bm_row[cb >> 5]
                                                  : mask_word32(i, cb, n, lrow, p.si, p.sl, p.mode);
#pragma unroll
          for (int e = 0; e < 32; ++e) xr[e] = ((bits >> e) & 1u) ? xr[e] : 0xff800000u;
        }
      };
      // rare: refresh the shared reference max from the raw scores of BOTH halves, rescale this half's O columns
      auto refresh = [&]() {
        ptx::mbar_wait(bar(bSFull + g), j & 1);               // H1(j) complete => PV(j-1) complete (commit order)
        ptx::tc_fence_after();
        if (h == 0) wait_block();
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t xr[32];
          ptx::tmem_ld32(tB + (uint32_t)(c * 32), xr);
          ptx::tc_wait_ld();
          masked(c, xr);
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            mx0 = fmaxf(mx0, fmaxf(__uint_as_float(xr[e]), __uint_as_float(xr[e + 1])));
            mx1 = fmaxf(mx1, fmaxf(__uint_as_float(xr[e + 2]), __uint_as_float(xr[e + 3])));
          }
        }
        const float mine = fmaxf(mx0, mx1);
        const float m_new = fmaxf(m_ref, fmaxf(mine, exchange(mine)) * sc);
        if (j > 0) {
          const float alpha = (m_new == -INFINITY) ? 1.0f : fast_exp2(m_ref - m_new);   // m_ref = -inf -> 0
          l_run *= alpha;
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            uint32_t r[32];
            ptx::tmem_ld32(tO + c * 32, r);
            ptx::tc_wait_ld();
#pragma unroll
            for (int e = 0; e < 32; ++e) r[e] = __float_as_uint(__uint_as_float(r[e]) * alpha);
            ptx::tmem_st32(tO + c * 32, r);
          }
          ptx::tc_wait_st();
        }
        m_ref = m_new;
      };
      if (__any_sync(0xffffffffu, m_ref == -INFINITY)) refresh();
      // optimistic pass over this half's 64 scores (no row max; see the 8-warp pass below for the criterion)
      uint32_t pk[32];             // this half's 64 probabilities as bf16 pairs
      uint64_t sum2;
#pragma unroll 1
      for (int attempt = 0;; ++attempt) {
        const float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
        const uint64_t negm2 = pk_f2(neg_m, neg_m);
        sum2 = pk_f2(0.f, 0.f);
        auto phase_a = [&](int c, uint32_t (&xr)[32]) {       // raw scores -> t = x * scale - m_ref, in place
          masked(c, xr);
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            float a, b;
            upk_f2(ffma2(pk_f2(__uint_as_float(xr[e]), __uint_as_float(xr[e + 1])), sc2, negm2), a, b);
            xr[e] = __float_as_uint(a);
            xr[e + 1] = __float_as_uint(b);
          }
        };
        auto phase_b = [&](int c, const uint32_t (&xr)[32]) { // t -> probabilities, row sum, packed bf16 pairs
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            float a = __uint_as_float(xr[e]), b = __uint_as_float(xr[e + 1]);
            uint64_t q2;
            if ((e & 15) < kEmu) {
              q2 = exp2_poly2(pk_f2(a, b));
              upk_f2(q2, a, b);
            } else {
              a = fast_exp2(a);
              b = fast_exp2(b);
              q2 = pk_f2(a, b);
            }
            sum2 = fadd2(sum2, q2);
            pk[c * 16 + e / 2] = pack2(a, b);
          }
        };
        wait_block();
        uint32_t b0[32], b1[32];
        ptx::tmem_ld32(tB, b0);
        ptx::tc_wait_ld();
        ptx::tmem_ld32(tB + 32, b1);
        phase_a(0, b0);
        ptx::tc_wait_ld();
        phase_a(1, b1);
        phase_b(0, b0);
        float stale = 0.f;
        if (attempt == 0) {
          float sa, sb, t0 = -INFINITY, t1 = -INFINITY;
          upk_f2(sum2, sa, sb);
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            t0 = fmaxf(t0, fmaxf(__uint_as_float(b1[e]), __uint_as_float(b1[e + 1])));
            t1 = fmaxf(t1, fmaxf(__uint_as_float(b1[e + 2]), __uint_as_float(b1[e + 3])));
          }
          stale = (!(sa + sb <= 128.0f) || fmaxf(t0, t1) > 8.0f) ? 1.f : 0.f;
          // both threads of a row must take the same decision: every score of the tile is in registers by now
          stale = fmaxf(stale, exchange(stale));
        }
        if (!__any_sync(0xffffffffu, stale != 0.f)) {
          if (h == 1) {           // H1(j)'s block is read and will not be re-read: the leader may issue H0(j+1) into it
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_leader(bar(bH1Read + g));
          }
          phase_b(1, b1);
          break;
        }
        refresh();
      }
      {
        float a, b;
        upk_f2(sum2, a, b);
        l_run += a + b;
      }
      // bf16 P: keys 64h .. 64h+63 of the tile -> columns [32h, 32h + 32) of H0's block (the other half loaded its
      // scores from there before the exchange barrier)
      ptx::tmem_st32(tH0 + (uint32_t)(h * 32), pk);
      ptx::tc_wait_st();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_leader(bar(bPReady + g));
    }
    // ---- epilogue: this half's 64 columns of O / l -> bf16 -> global ------------------------------
    ptx::mbar_wait(bar(bOFull + g), 0);
    ptx::tc_fence_after();
    const float l_tot = l_run + exchange(l_run);
    const float inv = l_tot > 0.f ? fast_rcp(l_tot) : 0.f;
    __nv_bfloat16* orow = p.out + (long long)(seq_start + i) * p.o_stride + head * kD + h * 64;   // seq_start + i = query row
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t r[32];
      ptx::tmem_ld32(tO + c * 32, r);
      ptx::tc_wait_ld();
      if (i <= r1) {
#pragma unroll
        for (int v4 = 0; v4 < 4; ++v4) {
          uint4 o;
          o.x = pack2(__uint_as_float(r[v4 * 8 + 0]) * inv, __uint_as_float(r[v4 * 8 + 1]) * inv);
          o.y = pack2(__uint_as_float(r[v4 * 8 + 2]) * inv, __uint_as_float(r[v4 * 8 + 3]) * inv);
          o.z = pack2(__uint_as_float(r[v4 * 8 + 4]) * inv, __uint_as_float(r[v4 * 8 + 5]) * inv);
          o.w = pack2(__uint_as_float(r[v4 * 8 + 6]) * inv, __uint_as_float(r[v4 * 8 + 7]) * inv);
          *reinterpret_cast<uint4*>(orow + c * 32 + v4 * 8) = o;
        }
      }
    }
    if (p.lse != nullptr && h == 0 && i <= r1)
      p.lse[(long long)head * p.total_T + seq_start + i] = m_ref * 0.6931471805599453f + logf(l_tot);
    ptx::tc_fence_before();
  } else if (!kSplit && warp >= 8) {
    // (the second issuer warp of a variant with ONE issuer, and both issuer warps of the peer CTA: nothing to do)
  } else {
    // =========================== softmax groups (both CTAs) ===========================
    const int g = warp >> 2;                       // 0: head h0, 1: head h0 + 1
    const int row_in_tile = (warp & 3) * 32 + lane;
    const int i = r0 + row_in_tile;                // query row inside the sequence
    const int head = h0 + g;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + lane_base + kColS + (uint32_t)(g * 128);
    const uint32_t tO = tmem_base + lane_base + kColO + (uint32_t)(g * 128);
    const float sc = p.scale_log2;                 // > 0 (checked by the launcher)
    float m_ref = -INFINITY, l_run = 0.f;          // running max (scaled, log2 domain) and row sum
    const uint32_t* bm_row = p.bitmask ? p.bitmask + (long long)min(i, n - 1) * p.bitmask_stride : nullptr;
    const uint64_t sc2 = pk_f2(sc, sc);

    for (int j = 0; j < total; ++j) {
      if constexpr (kAhead) ptx::mbar_wait(bar(bH0Full + g), j & 1);          // first 64 keys of S(j)
      else ptx::mbar_wait(bar(bSFull + g), j & 1);
      if (warp < 4) { PAIR_TRACE(warp, j); }
      ptx::tc_fence_after();
      const int c0 = tile_of(j) * kBN;
      const bool full = bm_row == nullptr &&
                        tile_is_full(r0, r1, c0, c0 + kBN - 1, n, lrow, p.si, p.sl, p.mode);
      uint32_t pk[64];             // the row's 128 probabilities as bf16 pairs
      uint64_t sum2;               // two partial row sums
      if constexpr (kAhead) {
        // column blocks of this tile's two 64-key halves (they swap every tile); P goes to H0's
        const uint32_t tH0 = tS + (uint32_t)((j & 1) * 64);
        const uint32_t tH1 = tS + (uint32_t)(((j & 1) ^ 1) * 64);
        auto masked = [&](int c, uint32_t (&xr)[32]) {          // apply the mask of chunk c (PARTIAL tiles only)
          if (!full) {
            const int cb = pin_in_branch(c0 + c * 32);
            const uint32_t bits = bm_row != nullptr ? bm_row[cb >> 5]
                                                    : mask_word32(i, cb, n, lrow, p.si, p.sl, p.mode);
#pragma unroll
            for (int e = 0; e < 32; ++e) xr[e] = ((bits >> e) & 1u) ? xr[e] : 0xff800000u;
          }
        };
        // Refresh of the reference max: a separate sweep over the tile's 128 raw scores (still intact in
        // TMEM: P is stored last), then O and l are rescaled.  Runs on the first tile, while a row has seen
        // no unmasked score yet, and when the optimistic pass below overflows -- rarely.
        auto refresh = [&]() {
          ptx::mbar_wait(bar(bSFull + g), j & 1);
          ptx::tc_fence_after();
          float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t xr[32];
            ptx::tmem_ld32((c < 2 ? tH0 : tH1) + (uint32_t)((c & 1) * 32), xr);
            ptx::tc_wait_ld();
            masked(c, xr);
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
              mx0 = fmaxf(mx0, fmaxf(__uint_as_float(xr[e]), __uint_as_float(xr[e + 1])));
              mx1 = fmaxf(mx1, fmaxf(__uint_as_float(xr[e + 2]), __uint_as_float(xr[e + 3])));
            }
          }
          const float m_new = fmaxf(m_ref, fmaxf(mx0, mx1) * sc);
          if (j > 0) {
            // S(j) complete => PV(j-1) complete (commit order); PV(j) waits for our arrive below
            const float alpha = (m_new == -INFINITY) ? 1.0f : fast_exp2(m_ref - m_new);   // m_ref = -inf -> 0
            l_run *= alpha;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint32_t r[32];
              ptx::tmem_ld32(tO + c * 32, r);
              ptx::tc_wait_ld();
#pragma unroll
              for (int e = 0; e < 32; ++e) r[e] = __float_as_uint(__uint_as_float(r[e]) * alpha);
              ptx::tmem_st32(tO + c * 32, r);
            }
            ptx::tc_wait_st();
          }
          m_ref = m_new;
        };
        if (__builtin_expect(__any_sync(0xffffffffu, m_ref == -INFINITY), 0)) refresh();
        // The optimistic pass: exponentiate against the CURRENT reference without looking for the row max at
        // all (96 FMNMX per row and tile saved).  A stale reference shows up as large probabilities: the pass
        // stands while the row sum of chunks 0-2 stays <= 2^8 and no score of chunk 3 exceeds the reference by
        // 2^8 (checked before P is stored).
        // H1(j)'s column block is handed back to the MMA issuer as soon as its 64 scores sit in registers --
        // before they are exponentiated and before the stale check (ncu round 2: the softmax warps spent 17 % of
        // their time waiting for H0 of the NEXT tile, whose issue used to wait for that check).  A stale pass
        // therefore never re-reads H1 from TMEM: late_refresh() takes chunks 2-3 from registers (t values,
        // re-based) and re-reads only H0's block, which this warp owns until it stores P there.
        float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
        uint64_t negm2 = pk_f2(neg_m, neg_m);
        sum2 = pk_f2(0.f, 0.f);
        // A: raw scores of chunk c (in xr) -> t = x * scale - m_ref, in place
        auto phase_a = [&](int c, uint32_t (&xr)[32]) {
          masked(c, xr);
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            float a, b;
            upk_f2(ffma2(pk_f2(__uint_as_float(xr[e]), __uint_as_float(xr[e + 1])), sc2, negm2), a, b);
            xr[e] = __float_as_uint(a);
            xr[e + 1] = __float_as_uint(b);
          }
        };
        // B: t of chunk c (in xr) -> probabilities, row sum, 16 packed bf16 pairs
        auto phase_b = [&](int c, const uint32_t (&xr)[32]) {
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            float a = __uint_as_float(xr[e]), b = __uint_as_float(xr[e + 1]);
            uint64_t q2;
            if ((((e >> 1) * (kEmu / 2)) & 7) < kEmu / 2) {        // kEmu of every 16, spread evenly over the pairs
              q2 = exp2_poly2(pk_f2(a, b));
              upk_f2(q2, a, b);
            } else {
              a = fast_exp2(a);
              b = fast_exp2(b);
              q2 = pk_f2(a, b);
            }
            sum2 = fadd2(sum2, q2);
            pk[c * 16 + e / 2] = pack2(a, b);
          }
        };
        uint32_t b0[32], b1[32], b2[32];
        ptx::tmem_ld32(tH0, b0);
        ptx::tc_wait_ld();
        ptx::tmem_ld32(tH0 + 32, b1);
        phase_a(0, b0);
        ptx::tc_wait_ld();
        // everything that only needs H0 first (half of the pass), then wait for keys 64..127
        phase_a(1, b1);
        phase_b(0, b0);
        phase_b(1, b1);
        ptx::mbar_wait(bar(bSFull + g), j & 1);   // (completed long ago after a refresh)
        ptx::tc_fence_after();
        ptx::tmem_ld32(tH1, b2);
        ptx::tc_wait_ld();
        ptx::tmem_ld32(tH1 + 32, b0);
        phase_a(2, b2);
        ptx::tc_wait_ld();
        // H1(j)'s block is read and will not be re-read: the leader may issue H0(j+1) into it
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive_leader(bar(bH1Read + g));
        phase_a(3, b0);
        phase_b(2, b2);
        bool stale;
        {
          float sa, sb, t0 = -INFINITY, t1 = -INFINITY;
          upk_f2(sum2, sa, sb);
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            t0 = fmaxf(t0, fmaxf(__uint_as_float(b0[e]), __uint_as_float(b0[e + 1])));
            t1 = fmaxf(t1, fmaxf(__uint_as_float(b0[e + 2]), __uint_as_float(b0[e + 3])));
          }
          stale = !(sa + sb <= 256.0f) || fmaxf(t0, t1) > 8.0f;        // (!(<=) also catches NaN / inf)
        }
        if (__builtin_expect(__any_sync(0xffffffffu, stale), 0)) {
          // late refresh: true row max of this tile = max(raw H0 scores from TMEM, t of chunks 2-3 + old reference)
          float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            mx0 = fmaxf(mx0, fmaxf(__uint_as_float(b2[e]), __uint_as_float(b2[e + 1])));
            mx1 = fmaxf(mx1, fmaxf(__uint_as_float(b0[e]), __uint_as_float(b0[e + 1])));
          }
          float m_new = fmaxf(m_ref, fmaxf(mx0, mx1) - neg_m);      // t = x * scale + neg_m
#pragma unroll 1
          for (int c = 0; c < 2; ++c) {
            ptx::tmem_ld32(tH0 + (uint32_t)(c * 32), b1);
            ptx::tc_wait_ld();
            masked(c, b1);
            float r0_ = -INFINITY, r1_ = -INFINITY;
#pragma unroll
            for (int e = 0; e < 32; e += 2) {
              r0_ = fmaxf(r0_, __uint_as_float(b1[e]));
              r1_ = fmaxf(r1_, __uint_as_float(b1[e + 1]));
            }
            m_new = fmaxf(m_new, fmaxf(r0_, r1_) * sc);
          }
          if (j > 0) {
            // S(j) complete => PV(j-1) complete (commit order); PV(j) waits for our "P ready" below
            const float alpha = (m_new == -INFINITY) ? 1.0f : fast_exp2(m_ref - m_new);   // m_ref = -inf -> 0
            l_run *= alpha;
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
              ptx::tmem_ld32(tO + c * 32, b1);
              ptx::tc_wait_ld();
#pragma unroll
              for (int e = 0; e < 32; ++e) b1[e] = __float_as_uint(__uint_as_float(b1[e]) * alpha);
              ptx::tmem_st32(tO + c * 32, b1);
            }
            ptx::tc_wait_st();
          }
          m_ref = m_new;
          const float neg_new = (m_ref == -INFINITY) ? 0.f : -m_ref;
          const float delta = neg_new - neg_m;                       // re-base the t values still in registers
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            b2[e] = __float_as_uint(__uint_as_float(b2[e]) + delta);
            b0[e] = __float_as_uint(__uint_as_float(b0[e]) + delta);
          }
          neg_m = neg_new;
          negm2 = pk_f2(neg_m, neg_m);
          sum2 = pk_f2(0.f, 0.f);
          // redo chunks 0-1 from TMEM (one at a time through b1) and chunk 2 from registers
          ptx::tmem_ld32(tH0, b1);
          ptx::tc_wait_ld();
          phase_a(0, b1);
          phase_b(0, b1);
          ptx::tmem_ld32(tH0 + 32, b1);
          ptx::tc_wait_ld();
          phase_a(1, b1);
          phase_b(1, b1);
          phase_b(2, b2);
        }
        // the first half of P can go out while the last chunk is still being exponentiated
        ptx::tmem_st32(tH0, *reinterpret_cast<uint32_t(*)[32]>(&pk[0]));
        phase_b(3, b0);
        {
          float a, b;
          upk_f2(sum2, a, b);
          l_run += a + b;
        }
        ptx::tmem_st32(tH0 + 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[32]));
        ptx::tc_wait_st();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive_leader(bar(bPReady + g));
        if (warp < 4) { PAIR_TRACE(4 + warp, j); }
        continue;
      }
#pragma unroll 1
      for (int attempt = 0;; ++attempt) {
        const float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
        const uint64_t negm2 = pk_f2(neg_m, neg_m);
        float mx0 = -INFINITY, mx1 = -INFINITY;
        sum2 = pk_f2(0.f, 0.f);
        // 32 scores (columns c0 + 32c ...) -> 16 packed bf16 pairs
        auto chunk = [&](int c, uint32_t (&xr)[32]) {
          float x[32];
#pragma unroll
          for (int e = 0; e < 32; ++e) x[e] = __uint_as_float(xr[e]);
          if (!full) {
            const int cb = pin_in_branch(c0 + c * 32);
            const uint32_t bits = bm_row != nullptr ? bm_row[cb >> 5]
                                                    : mask_word32(i, cb, n, lrow, p.si, p.sl, p.mode);
#pragma unroll
            for (int e = 0; e < 32; ++e) x[e] = ((bits >> e) & 1u) ? x[e] : -INFINITY;
          }
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            mx0 = fmaxf(mx0, fmaxf(x[e], x[e + 1]));
            mx1 = fmaxf(mx1, fmaxf(x[e + 2], x[e + 3]));
          }
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            uint64_t q2 = ffma2(pk_f2(x[e], x[e + 1]), sc2, negm2);
            float a, b;
            if ((e & 15) < kEmu) {
              q2 = exp2_poly2(q2);
              upk_f2(q2, a, b);
            } else {
              upk_f2(q2, a, b);
              a = fast_exp2(a);
              b = fast_exp2(b);
              q2 = pk_f2(a, b);
            }
            sum2 = fadd2(sum2, q2);
            pk[c * 16 + e / 2] = pack2(a, b);
          }
        };
        uint32_t xa[32], xb[32];
        ptx::tmem_ld32(tS, xa);
        ptx::tc_wait_ld();
        ptx::tmem_ld32(tS + 32, xb);        // in flight while chunk 0 is processed
        chunk(0, xa);
        ptx::tc_wait_ld();
        ptx::tmem_ld32(tS + 64, xa);
        chunk(1, xb);
        ptx::tc_wait_ld();
        ptx::tmem_ld32(tS + 96, xb);
        chunk(2, xa);
        ptx::tc_wait_ld();
        chunk(3, xb);
        const float m_new = fmaxf(m_ref, fmaxf(mx0, mx1) * sc);
        const bool grow = (m_new > m_ref + 8.0f) || (m_ref == -INFINITY && m_new > -INFINITY);
        if (attempt == 1 || !__any_sync(0xffffffffu, grow)) break;
        if (j > 0) {
          // S(j) complete => PV(j-1) complete (commit order); PV(j) waits for our arrive below
          const float alpha = (m_new == -INFINITY) ? 1.0f : fast_exp2(m_ref - m_new);   // m_ref=-inf -> 0
          l_run *= alpha;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t r[32];
            ptx::tmem_ld32(tO + c * 32, r);
            ptx::tc_wait_ld();
#pragma unroll
            for (int e = 0; e < 32; ++e) r[e] = __float_as_uint(__uint_as_float(r[e]) * alpha);
            ptx::tmem_st32(tO + c * 32, r);
          }
          ptx::tc_wait_st();
        }
        m_ref = m_new;
      }
      {
        float a, b;
        upk_f2(sum2, a, b);
        l_run += a + b;
      }
      // bf16 P into columns [0,64) of S (every score of the tile has been read by now)
      ptx::tmem_st32(tS, *reinterpret_cast<uint32_t(*)[32]>(&pk[0]));
      ptx::tmem_st32(tS + 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[32]));
      ptx::tc_wait_st();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_leader(bar(bPReady + g));
    }
    // ---- epilogue: O / l -> bf16 -> global ------------------------------------
    ptx::mbar_wait(bar(bOFull + g), 0);
    ptx::tc_fence_after();
    const float inv = l_run > 0.f ? fast_rcp(l_run) : 0.f;
    __nv_bfloat16* orow = p.out + (long long)(seq_start + i) * p.o_stride + head * kD;   // seq_start + i = query row
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t r[32];
      ptx::tmem_ld32(tO + c * 32, r);
      ptx::tc_wait_ld();
      if (i <= r1) {
#pragma unroll
        for (int v4 = 0; v4 < 4; ++v4) {
          uint4 o;
          o.x = pack2(__uint_as_float(r[v4 * 8 + 0]) * inv, __uint_as_float(r[v4 * 8 + 1]) * inv);
          o.y = pack2(__uint_as_float(r[v4 * 8 + 2]) * inv, __uint_as_float(r[v4 * 8 + 3]) * inv);
          o.z = pack2(__uint_as_float(r[v4 * 8 + 4]) * inv, __uint_as_float(r[v4 * 8 + 5]) * inv);
          o.w = pack2(__uint_as_float(r[v4 * 8 + 6]) * inv, __uint_as_float(r[v4 * 8 + 7]) * inv);
          *reinterpret_cast<uint4*>(orow + c * 32 + v4 * 8) = o;
        }
      }
    }
    if (p.lse != nullptr && i <= r1)
      p.lse[(long long)head * p.total_T + seq_start + i] = m_ref * 0.6931471805599453f + logf(l_run);
    ptx::tc_fence_before();
  }

  // ---- teardown: neither CTA may leave (or free TMEM) while its peer still reads its memory ----------
  ptx::cluster_sync();
  if (warp == kWMma) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_pair(tmem_base, 512);
  }
}

}  // namespace

bool pair_supported(const AttnParams& p) {
  return p.Hkv > 0 && p.Hq % p.Hkv == 0 && (p.Hq / p.Hkv) % 4 == 0 && p.scale_log2 > 0.f && p.block_table == nullptr;
}

int launch_attn_sm100_pair(const AttnParams& p, cudaStream_t s) {
  if (!pair_supported(p)) return TMX_ERR_UNSUPPORTED;
  // Product: the look-ahead schedule with 4 of every 16 exponentials on the FMA pipe (measured best under the
  // power cap at 128K: 1178 vs 1155 TFLOP/s with all of them on MUFU; profiles/r02_pair_kernel.md).
  using Kern = void (*)(const AttnParams, const PairMaps);
  Kern kern = attn_sm100_pair_kernel<4, true>;
  int slot = kAttrV3Pair;
  bool ahead = true;
  int threads = kThreads;
#ifdef TMX_DEV
  // TMX_SM100_VARIANT (developer build): bits 0-1 = exponentials on the FMA pipe per 16 (0, 4, 6, 2 look-ahead / 8 classic),
  // bit 2 = classic schedule (one N = 128 score MMA per tile) instead of look-ahead
  {
    const int v = sm100_variant();
    static const Kern ahead_k[4] = {attn_sm100_pair_kernel<0, true>, attn_sm100_pair_kernel<4, true>,
                                    attn_sm100_pair_kernel<6, true>, attn_sm100_pair_kernel<2, true>};
    static const Kern classic_k[4] = {attn_sm100_pair_kernel<0, false>, attn_sm100_pair_kernel<4, false>,
                                      attn_sm100_pair_kernel<6, false>, attn_sm100_pair_kernel<8, false>};
    if (v != 0) { ahead = (v & 4) == 0; kern = ahead ? ahead_k[v & 3] : classic_k[v & 3]; slot = -1; }
    // bit 3: two threads per row (16 softmax warps); bits 0-1 pick 0 / 4 / 6 / 2 emulated exponentials per 16
    if (v & 8) {
      static const Kern split_k[4] = {attn_sm100_pair_kernel<0, true, true>, attn_sm100_pair_kernel<4, true, true>,
                                      attn_sm100_pair_kernel<6, true, true>, attn_sm100_pair_kernel<2, true, true>};
      kern = split_k[v & 3]; ahead = true; threads = kThreadsSplit; slot = -1;
    }
  }
#endif
  if (slot < 0) TMX_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  else TMX_SET_SMEM_ATTR(kern, slot, kSmem);
  PairMaps maps;
  int st = tma_map_2d(&maps.q, p.q, p.total_T, (long long)p.Hq * kD, p.q_stride, kBM);
  if (st == TMX_OK) st = tma_map_2d(&maps.k, p.k, p.total_kv, (long long)p.Hkv * kD, p.kv_stride, ahead ? kBN / 4 : kBN / 2);
  if (st == TMX_OK) st = tma_map_2d(&maps.v, p.v, p.total_kv, (long long)p.Hkv * kD, p.kv_stride, kBN);
  if (st != TMX_OK) return st;
  dim3 grid((unsigned)(2 * p.n_slots), (unsigned)(p.Hq / 4));     // clusters of 2 along x
  kern<<<grid, threads, kSmem, s>>>(p, maps);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

}  // namespace tmx

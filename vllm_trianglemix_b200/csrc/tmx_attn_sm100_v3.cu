// Masked GQA prefill attention on the 5th-generation tensor cores, 128-key tiles (kernel v3).
//
// One CTA = one 128-token query tile x TWO query heads that share a KV head (GQA reuse:
// every K/V tile is fetched once for 256 query rows).  10 warps:
//   warps 0-3  softmax group A (head h0):   one thread per query row, TMEM lanes 0..127
//   warps 4-7  softmax group B (head h0+1)
//   warp  8    TMA producer: Q once, then K(0) V(0) K(1) V(1) ... of the tiles the block index
//              lists, through ONE ring of 5 x 32 KB stages (consumption order = load order)
//   warp  9    MMA issuer (warp-uniform loop, one elected lane issues tcgen05.mma)
// KV tiles are 128 keys, so both contractions run as M128 N128 MMAs (the shape that reaches
// the tensor-pipe peak; the 64-key kernel's N=64 score MMAs top out ~15 % lower).
// TMEM (512 columns): S_A | S_B (128 fp32 columns each) | O_A | O_B (128 each).  bf16 P
// overwrites columns [0,64) of the S buffer it came from and is read by the MMA from TMEM.
// Issue order   S_A(0) S_B(0) { PV_A(j) S_A(j+1) PV_B(j) S_B(j+1) }_j :
// while one group exponentiates, the tensor pipe runs the other group's PV and next scores.
// "S_g(j+1) complete" implies "PV_g(j) complete" (commit order), so a group may rescale its
// O accumulator in TMEM right after it sees its next scores -- done lazily, only when some
// row's running max grew by more than 2^8.  A share of the exponentials runs on the FMA pipe
// (cubic 2^f, Cody-Waite split) because MUFU.EX2 at 16/clk/SM costs exactly as many cycles
// per tile as the two MMAs.  Fully masked KV tiles are never loaded; only PARTIAL tiles
// evaluate the mask predicate (or read the bitmask).
//
// Replaces vllm/attention/ascend_trianglemix.py:86-106 of the reference for dense-causal
// layers (and for any triangle setting through the same block index).
#include <cstdlib>
#include "tmx_attn.cuh"
#include "tmx_sm100_ptx.cuh"
#include "tmx_tma.cuh"

namespace tmx {
namespace {

constexpr int kD = 128;
constexpr int kBM = 128;
constexpr int kBN = 128;
constexpr int kThreads = 320;
constexpr int kQBytes = 2 * 2 * kBM * 128;        // 2 heads x 2 panels of [128 rows][64 d] = 64 KB
constexpr int kPanelBytes = kBN * 128;            // [128 keys][64 d] bf16 = 16 KB
constexpr int kTileBytes = 2 * kPanelBytes;       // K or V tile = 32 KB
constexpr int kStages = 5;
constexpr int kOffRing = kQBytes;
constexpr int kOffBar = kOffRing + kStages * kTileBytes;

constexpr int bQFull = 0;
constexpr int bFull = 1;                          // [kStages]
constexpr int bEmpty = bFull + kStages;           // [kStages]
constexpr int bSFull = bEmpty + kStages;          // [2 groups]
constexpr int bPReady = bSFull + 2;               // [2 groups], 4 arrivals (one per warp)
constexpr int bOFull = bPReady + 2;               // [2 groups]
constexpr int bP1Ready = bOFull + 2;              // [2 groups] kSplitP: second half of P stored
constexpr int bPV0Done = bP1Ready + 2;            // [2 groups] kSplitP: first half of PV(j) complete
constexpr int kNumBars = bPV0Done + 2;
constexpr int kSmem = kOffBar + kNumBars * 8 + 16 + 1024;   // + alignment slack
static_assert(kSmem <= 232448, "exceeds the 227 KB of shared memory a CTA may use");

constexpr uint32_t kColS = 0;      // + g * 128
constexpr uint32_t kColO = 256;    // + g * 128

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// 2^t on the FMA/ALU pipes (no MUFU): round-to-nearest split t = n + f, f in [-0.5, 0.5],
// cubic minimax for 2^f (max rel. error 7.5e-5, far below bf16's 2^-9), n added into the
// exponent field.  t is clamped at -126 (result then ~1e-38, i.e. 0 for bf16).
__device__ __forceinline__ float exp2_poly(float t) {
  t = fmaxf(t, -126.0f);
  const float r = t + 12582912.0f;          // 1.5 * 2^23: low mantissa bits of r hold n
  const float f = t - (r - 12582912.0f);
  float pl = fmaf(0.0551716685f, f, 0.2426111251f);
  pl = fmaf(pl, f, 0.6932609677f);
  pl = fmaf(pl, f, 0.9999280572f);
  return __int_as_float(__float_as_int(pl) + (__float_as_int(r) << 23));
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
// exp2_poly on a pair
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
// kDbg (developer timing experiments, results are garbage): 1 = MMA/TMA side alone,
// 2 = softmax side alone (S tile 0 computed once and re-read), 3 = group A alone (serial chain).
//
// kTail: split-KV worker for the dense tail rows of a TRIANGLE layer (the `last - 1` rows at the
// end of every sequence attend all earlier keys: 2/3 of a 128K triangle layer's FLOPs sit in
// its last 63 query rows).  The CTA then takes ALL G = 2P query heads of one KV head for a
// tile of TT = 128 / P tokens -- packed row m of group g = token m / P, head h0 + g P + m % P,
// loaded by a 3-D TMA box [64 d][P heads][TT tokens] that lands in the same K-major layout --
// and one contiguous chunk of the KV tiles, and writes an fp32 partial (O, m, l) that
// combine_tail_kernel merges.  K/V are therefore read once per KV head (not once per query
// head) and the work spreads over ~148 CTAs instead of one.
//
// kSplitP: P is stored in two halves (keys [0,64) and [64,128) of the tile); the MMA warp issues
// the first four k-steps of PV as soon as the first half is there, so that only HALF a PV (plus
// the next score MMA) sits between the end of a group's exponentials and its next S tile.
//
// kSm == 2 (kPipe): the softmax pass software-pipelined across its four 32-column chunks.  In the
// plain pass each chunk is a scheduling island between two tcgen05.wait::ld: ptxas interleaves its
// MUFU.EX2 with the chunk's FMA-pipe work until that work runs out, and the last ~12 of the 24
// MUFUs of every chunk then issue back to back at 8 clk each with nothing to hide behind (seen in
// the SASS).  Here a chunk is split into A (mask, row max, t = x*scale - m) and B (exponentials,
// row sum, bf16 pack), and each region between two waits holds B(c) next to A(c+1), through three
// rotating 32-register score buffers.
//
// kSm == 3 (kAhead): kPipe + the score MMA of a tile issued as two 64-key halves H0 | H1 living in
// the two 64-column blocks of the group's S buffer, the blocks swapping roles every tile:
//   tile j:  H0(j) in block j&1, H1(j) in the other; P(j) (64 columns) overwrites H0(j)'s block.
// The block of H1(j) is free as soon as the softmax pass has loaded it and knows it will not redo
// the pass (row max is known after phase A of the last chunk, ~60 % into the pass), so H0(j+1) is
// issued THEN -- before P(j) exists -- and H1(j+1) follows PV(j) into P(j)'s block.  The next
// pass starts on H0(j+1) right after its P store and needs H1(j+1) only half-way through, which
// takes most of PV(j) + S(j+1) out of the per-group chain softmax -> P -> PV -> S -> softmax.
// Barriers: bPV0Done = "H0 ready", bSFull = "H1 ready", bP1Ready = "H1's block free".
template <int kEmu, int kDbg, bool kTail, int kSm>
__global__ void __launch_bounds__(kThreads, 1)
attn_sm100_v3_kernel(const AttnParams p, const __grid_constant__ TmaMaps maps) {
  constexpr bool kSplitP = (kSm == 1);
  constexpr bool kAhead = (kSm == 3) || (kSm == 4);
  constexpr bool kPoll = (kSm == 4);   // kAhead with an event-driven MMA issuer (each group advances on its own)
  constexpr bool kPipe = (kSm == 2) || kAhead;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;   // SWIZZLE_128B: 1024-byte atoms
  const uint32_t sQ = smem_base, sRing = smem_base + kOffRing;
  const uint32_t sBar = smem_base + kOffBar;
  const uint32_t sTmemPtr = sBar + kNumBars * 8;
  auto bar = [&](int i) { return sBar + 8u * (uint32_t)i; };

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // provably warp-uniform
  const int lane = threadIdx.x & 31;

  // ---- work item (uniform per CTA) -------------------------------------------
  // r0 / r1 / i are ABSOLUTE rows of the (context + chunk) sequence; q_row0 is the row of the
  // query buffer that holds absolute row r0 (they differ by the sequence's context length).
  int n, r0, r1, lo0, cnt0, lo1, total, seq_start, h0, kvh, lrow, seq = 0;
  int q_row0, k_start;
  int pshift = 0;          // log2(P): query heads packed per group (kTail only)
  long long part = -1;     // split-KV partial index (kTail only)
  if constexpr (kTail) {
    const int G = p.Hq / p.Hkv;
    pshift = (G >= 8) ? 2 : (G >= 4 ? 1 : 0);
    const int TT = kBM >> pshift;
    int e = blockIdx.x;
    const int chunk = e % p.nsplit_max; e /= p.nsplit_max;
    const int ql = e % p.last_tiles;
    seq = e / p.last_tiles;
    seq_start = p.cu[seq];
    n = p.cu[seq + 1] - seq_start;
    lrow = first_dense_row(n, p.last);
    if (n <= 0 || lrow >= n) return;
    r0 = lrow + ql * TT;           // tail tiles start AT the first dense row: every row is dense
    if (r0 >= n) return;
    r1 = min(n, r0 + TT) - 1;
    const int tiles = r1 / kBN + 1;
    const int ns = tail_nsplit(tiles, p.nsplit_max);
    if (chunk >= ns) return;
    lo0 = (int)((long long)chunk * tiles / ns);
    cnt0 = (int)((long long)(chunk + 1) * tiles / ns) - lo0;
    lo1 = 0; total = cnt0;
    kvh = blockIdx.y; h0 = kvh * G;
    part = (long long)(seq * p.last_tiles + ql) * p.nsplit_max + chunk;
    q_row0 = seq_start + r0; k_start = seq_start;
  } else {
    const SlotInfo slot = p.slots[p.n_slots - 1 - blockIdx.x];      // descending: longest first
    if (slot.seq < 0) return;
    seq = slot.seq;
    const SeqGeom gm = seq_geom(p, seq);
    n = slot.n; r0 = gm.ctx + slot.q_tile * kBM;
    r1 = min(n, r0 + kBM) - 1;
    lrow = first_dense_row(n, p.last);
    if (p.tail_clip && p.mode == TMX_MODE_TRIANGLE && lrow < n) r1 = min(r1, lrow - 1);
    lo0 = slot.lo0; cnt0 = slot.hi0 - slot.lo0; lo1 = slot.lo1;
    total = cnt0 + (slot.hi1 - slot.lo1);
    seq_start = gm.q_start - gm.ctx;      // query-buffer row of absolute row 0 (may be negative)
    q_row0 = gm.q_start + slot.q_tile * kBM; k_start = gm.k_start;
    h0 = 2 * blockIdx.y;
    kvh = h0 / (p.Hq / p.Hkv);
  }
  auto tile_of = [&](int it) { return it < cnt0 ? lo0 + it : lo1 + (it - cnt0); };

  // ---- setup -------------------------------------------------------------------
  if (warp == 8 && lane == 0) {
    ptx::prefetch_tmap(&maps.q); ptx::prefetch_tmap(&maps.k); ptx::prefetch_tmap(&maps.v);
    ptx::mbar_init(bar(bQFull), 1);
    for (int s = 0; s < kStages; ++s) { ptx::mbar_init(bar(bFull + s), 1); ptx::mbar_init(bar(bEmpty + s), 1); }
    for (int g = 0; g < 2; ++g) {
      ptx::mbar_init(bar(bSFull + g), 1);
      ptx::mbar_init(bar(bPReady + g), 4);       // one arrive per softmax warp of the group
      ptx::mbar_init(bar(bOFull + g), 1);
      ptx::mbar_init(bar(bP1Ready + g), 4);
      ptx::mbar_init(bar(bPV0Done + g), 1);
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
      ptx::mbar_expect_tx(bar(bQFull), kQBytes);
      for (int g = 0; g < 2; ++g)
        for (int pnl = 0; pnl < 2; ++pnl)
          if constexpr (kTail)
            ptx::tma_load_3d(sQ + (g * 2 + pnl) * (kBM * 128), &maps.q, bar(bQFull), pnl * 64,
                             h0 + (g << pshift), q_row0);
          else
            ptx::tma_load_2d(sQ + (g * 2 + pnl) * (kBM * 128), &maps.q, bar(bQFull),
                             (h0 + g) * kD + pnl * 64, q_row0);
      const int items = (kDbg == 2) ? 1 : 2 * total;     // item t: even = K(t/2), odd = V(t/2)
      int st = 0;
      // kTail streams K/V from HBM (no reuse across CTAs).  A CTA's own TMA boxes are 128-byte
      // pieces at a pitch of Hkv * 256 bytes; the Hkv CTAs of a chunk together cover whole rows.
      // Each of them therefore pulls its 1/Hkv share of the rows of a tile kPfDist tiles ahead
      // into L2 at FULL row width (contiguous Hkv * 256-byte bursts), so that DRAM sees long
      // sequential reads and the later box loads of all Hkv CTAs hit L2.
      constexpr int kPfDist = 4;
      const int pf_rows = (kBN + p.Hkv - 1) / p.Hkv;
      auto prefetch_tile = [&](int idx) {
        if (idx >= total) return;
        const int rbeg = kvh * pf_rows, rend = min(kBN, rbeg + pf_rows);
        const long long tok0 = (long long)k_start + (long long)tile_of(idx) * kBN;
        for (int r = rbeg; r < rend; ++r) {
          if (tok0 + r >= p.total_T) break;
          ptx::prefetch_l2(p.k + (tok0 + r) * p.kv_stride, (uint32_t)(p.Hkv * kD * 2));
          ptx::prefetch_l2(p.v + (tok0 + r) * p.kv_stride, (uint32_t)(p.Hkv * kD * 2));
        }
      };
      if constexpr (kTail)
        for (int d = 0; d < kPfDist; ++d) prefetch_tile(d);
      // Paged cache (chunked prefill / decode): a 128-key tile is gathered from its pages, one TMA box
      // per page (page_size < 128) or one box inside a page (page_size >= 128).  Boxes past the last
      // key re-read the last valid page (finite data; the scores there are masked anyway).
      const int ps = p.page_size;
      const int box_rows = (p.block_table != nullptr && ps < kBN) ? ps : kBN;
      const int nbox = kBN / box_rows;
      for (int t = 0; t < items; ++t) {
        if constexpr (kTail) { if ((t & 1) == 0) prefetch_tile((t >> 1) + kPfDist); }
        if (t >= kStages) ptx::mbar_wait(bar(bEmpty + st), ((t / kStages) - 1) & 1);
        ptx::mbar_expect_tx(bar(bFull + st), kTileBytes);
        const int key0 = tile_of(t >> 1) * kBN;
        const CUtensorMap* m = (t & 1) ? &maps.v : &maps.k;
        for (int bx = 0; bx < nbox; ++bx) {
          int row;
          if (p.block_table != nullptr) {
            const int key = min(key0 + bx * box_rows, n - 1);
            const int page = p.block_table[(long long)seq * p.bt_stride + key / ps];
            row = page * ps + (ps > kBN ? key0 % ps : 0);
          } else {
            row = k_start + key0;
          }
          for (int pnl = 0; pnl < 2; ++pnl)
            ptx::tma_load_2d(sRing + st * kTileBytes + pnl * kPanelBytes + bx * box_rows * 128, m,
                             bar(bFull + st), kvh * kD + pnl * 64, row);
        }
        if (++st == kStages) st = 0;
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // =========================== MMA issuer ===========================
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const int ntot = __shfl_sync(0xffffffffu, total, 0);
    constexpr uint32_t kIdescS = ptx::instr_desc_bf16(128, kBN, false);
    constexpr uint32_t kIdescO = ptx::instr_desc_bf16(128, 128, true);
    const uint64_t dQ = ptx::smem_desc(sQ, 16, 1024);
    const uint64_t dK = ptx::smem_desc(sRing, 16, 1024);             // K tiles: K-major (d contiguous)
    const uint64_t dV = ptx::smem_desc(sRing, kPanelBytes, 1024);    // V tiles: MN-major, LBO = d-panel pitch

    // S_g = Q_g K^T of the tile in ring stage `st`: 8 k-steps of 16 over d (two 64-d panels)
    auto issue_s = [&](int g, int st) {
      const uint32_t d_tmem = tb + kColS + (uint32_t)(g * 128);
      const uint64_t qa = dQ + (uint64_t)((g * 2 * (kBM * 128)) >> 4);
      const uint64_t kb = dK + (uint64_t)((st * kTileBytes) >> 4);
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const uint32_t qo = ((ks >> 2) * (kBM * 128) + (ks & 3) * 32) >> 4;
        const uint32_t ko = ((ks >> 2) * kPanelBytes + (ks & 3) * 32) >> 4;
        ptx::mma_ss(d_tmem, qa + qo, kb + ko, kIdescS, ks > 0);
      }
    };
    // O_g += P_g V: 8 k-steps of 16 keys; P (bf16) read from columns [0,64) of S_g
    auto issue_pv = [&](int g, int st, bool acc, int ks0 = 0, int ks1 = kBN / 16) {
      const uint32_t d_tmem = tb + kColO + (uint32_t)(g * 128);
      const uint32_t a_tmem = tb + kColS + (uint32_t)(g * 128);
      const uint64_t vb = dV + (uint64_t)((st * kTileBytes) >> 4);
#pragma unroll
      for (int ks = ks0; ks < ks1; ++ks)
        ptx::mma_ts(d_tmem, a_tmem + (uint32_t)(ks * 8), vb + (uint64_t)((ks * 2048) >> 4), kIdescO,
                    acc || ks > 0);
    };

    // kAhead: half h (64 keys) of S_g for tile j into column block (j & 1) ^ h
    constexpr uint32_t kIdescS64 = ptx::instr_desc_bf16(128, 64, false);
    auto issue_s_half = [&](int g, int st, int j, int h) {
      const uint32_t d_tmem = tb + kColS + (uint32_t)(g * 128) + (uint32_t)((((j & 1) ^ h)) * 64);
      const uint64_t qa = dQ + (uint64_t)((g * 2 * (kBM * 128)) >> 4);
      const uint64_t kb = dK + (uint64_t)((st * kTileBytes + h * 64 * 128) >> 4);   // keys 64h.. of the tile
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const uint32_t qo = ((ks >> 2) * (kBM * 128) + (ks & 3) * 32) >> 4;
        const uint32_t ko = ((ks >> 2) * kPanelBytes + (ks & 3) * 32) >> 4;
        ptx::mma_ss(d_tmem, qa + qo, kb + ko, kIdescS64, ks > 0);
      }
    };
    auto issue_pv_blk = [&](int g, int st, bool acc, int j) {     // P(j) sits in block j & 1
      const uint32_t d_tmem = tb + kColO + (uint32_t)(g * 128);
      const uint32_t a_tmem = tb + kColS + (uint32_t)(g * 128) + (uint32_t)((j & 1) * 64);
      const uint64_t vb = dV + (uint64_t)((st * kTileBytes) >> 4);
#pragma unroll
      for (int ks = 0; ks < kBN / 16; ++ks)
        ptx::mma_ts(d_tmem, a_tmem + (uint32_t)(ks * 8), vb + (uint64_t)((ks * 2048) >> 4), kIdescO,
                    acc || ks > 0);
    };

    int st = 0, ph = 0;                               // ring position of the next item to consume
    auto advance = [&] { if (++st == kStages) { st = 0; ph ^= 1; } };
    ptx::mbar_wait(bar(bQFull), 0);
    ptx::mbar_wait(bar(bFull + 0), 0);                // K(0)
    ptx::tc_fence_after();
    if constexpr (kAhead) {
      if (ptx::elect_one()) {
        for (int g = 0; g < 2; ++g) {
          issue_s_half(g, 0, 0, 0); ptx::mma_commit(bar(bPV0Done + g));
          issue_s_half(g, 0, 0, 1); ptx::mma_commit(bar(bSFull + g));
        }
        ptx::mma_commit(bar(bEmpty + 0));
      }
      __syncwarp();
      advance();
      if constexpr (kPoll) {
        // Ring item 0 = K(0); tile j: V(j) = item 1 + 2j, K(j+1) = item 2 + 2j.  Each group walks its own
        // tile counter; an item's stage is released by whichever group finishes with it second (a commit
        // covers every MMA this thread issued before it).
        int jg[2] = {0, 0}, stage[2] = {0, 0}, cons[2] = {1, 1};   // stage 0: H0 of tile j+1 pending, 1: PV of tile j
        auto test = [&](int b, int parity) {
          return __shfl_sync(0xffffffffu, (int)ptx::mbar_test(bar(b), (uint32_t)parity), 0) != 0;
        };
        auto item_full = [&](int item) { return test(bFull + item % kStages, (item / kStages) & 1); };
        uint32_t idle = 0;
        while (jg[0] < ntot || jg[1] < ntot) {
          bool progress = false;
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            const int j = jg[g];
            if (j >= ntot) continue;
            const bool more = j + 1 < ntot;
            const int vitem = 1 + 2 * j, kitem = 2 + 2 * j;
            if (more && stage[g] == 0) {
              if (test(bP1Ready + g, j & 1) && item_full(kitem)) {        // H1(j)'s block has been read
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                  issue_s_half(g, kitem % kStages, j + 1, 0);
                  ptx::mma_commit(bar(bPV0Done + g));
                }
                __syncwarp();
                stage[g] = 1;
                progress = true;
              }
            }
            if ((!more || stage[g] == 1) && test(bPReady + g, j & 1) && item_full(vitem)) {
              ptx::tc_fence_after();
              if (ptx::elect_one()) {
                issue_pv_blk(g, vitem % kStages, j > 0, j);
                if (!more) ptx::mma_commit(bar(bOFull + g));
                if (cons[g ^ 1] > vitem) ptx::mma_commit(bar(bEmpty + vitem % kStages));
                if (more) {
                  issue_s_half(g, kitem % kStages, j + 1, 1);             // into P(j)'s block, behind PV(j)
                  ptx::mma_commit(bar(bSFull + g));
                  if (cons[g ^ 1] > kitem) ptx::mma_commit(bar(bEmpty + kitem % kStages));
                }
              }
              __syncwarp();
              cons[g] = (more ? kitem : vitem) + 1;
              jg[g] = j + 1;
              stage[g] = 0;
              progress = true;
            }
          }
          if (progress) idle = 0;
          else { __nanosleep(32); if (++idle > (1u << 24)) __trap(); }
        }
      } else
      for (int j = 0; j < ntot; ++j) {
        const int vst = st, vph = ph;
        advance();
        const int kst = st, kph = ph;
        const bool more = j + 1 < ntot;
        if (more) advance();
        ptx::mbar_wait(bar(bFull + vst), vph);
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          if (more) {
            if (g == 0) ptx::mbar_wait(bar(bFull + kst), kph);
            ptx::mbar_wait(bar(bP1Ready + g), j & 1);           // H1(j)'s block has been read
            ptx::tc_fence_after();
            if (ptx::elect_one()) { issue_s_half(g, kst, j + 1, 0); ptx::mma_commit(bar(bPV0Done + g)); }
            __syncwarp();
          }
          ptx::mbar_wait(bar(bPReady + g), j & 1);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            issue_pv_blk(g, vst, j > 0, j);
            if (!more) ptx::mma_commit(bar(bOFull + g));
            if (g == 1) ptx::mma_commit(bar(bEmpty + vst));
            if (more) {
              issue_s_half(g, kst, j + 1, 1);                   // into P(j)'s block, behind PV(j)
              ptx::mma_commit(bar(bSFull + g));
              if (g == 1) ptx::mma_commit(bar(bEmpty + kst));
            }
          }
          __syncwarp();
        }
      }
    } else {
    if (ptx::elect_one()) {
      issue_s(0, 0); ptx::mma_commit(bar(bSFull + 0));
      if (kDbg != 3) { issue_s(1, 0); ptx::mma_commit(bar(bSFull + 1)); }
      if (kDbg != 2) ptx::mma_commit(bar(bEmpty + 0));
      if (kDbg == 2) { ptx::mma_commit(bar(bOFull + 0)); ptx::mma_commit(bar(bOFull + 1)); }
    }
    __syncwarp();
    advance();
    for (int j = 0; j < ntot && kDbg != 2; ++j) {
      const int vst = st, vph = ph;
      advance();
      const int kst = st, kph = ph;
      const bool more = j + 1 < ntot;
      if (more) advance();
      ptx::mbar_wait(bar(bFull + vst), vph);
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        if (kDbg == 3 && g == 1) {       // one group alone: only release the ring stages
          if (ptx::elect_one()) { ptx::mma_commit(bar(bEmpty + vst)); if (more) ptx::mma_commit(bar(bEmpty + kst)); }
          __syncwarp();
          continue;
        }
        if (kDbg != 1) ptx::mbar_wait(bar(bPReady + g), j & 1);
        ptx::tc_fence_after();
        if constexpr (kSplitP) {
          if (ptx::elect_one()) {
            issue_pv(g, vst, j > 0, 0, 4);
            ptx::mma_commit(bar(bPV0Done + g));
          }
          __syncwarp();
          if (kDbg != 1) ptx::mbar_wait(bar(bP1Ready + g), j & 1);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            issue_pv(g, vst, true, 4, 8);
            if (!more) ptx::mma_commit(bar(bOFull + g));
            if (g == 1) ptx::mma_commit(bar(bEmpty + vst));
          }
          __syncwarp();
        } else {
        if (ptx::elect_one()) {
          issue_pv(g, vst, j > 0);
          if (!more) ptx::mma_commit(bar(bOFull + g));
          if (g == 1) ptx::mma_commit(bar(bEmpty + vst));
        }
        __syncwarp();
        }
        if (more) {
          if (g == 0) { ptx::mbar_wait(bar(bFull + kst), kph); ptx::tc_fence_after(); }
          if (ptx::elect_one()) {
            issue_s(g, kst);
            ptx::mma_commit(bar(bSFull + g));
            if (g == 1) ptx::mma_commit(bar(bEmpty + kst));
          }
          __syncwarp();
        }
      }
    }
    }   // !kAhead
  } else {
    // =========================== softmax groups ===========================
    const int g = warp >> 2;                       // 0: head h0, 1: head h0 + 1
    const int row_in_tile = (warp & 3) * 32 + lane;
    // query row inside the sequence, and its head (kTail: P heads interleaved per token)
    const int i = kTail ? r0 + (row_in_tile >> pshift) : r0 + row_in_tile;
    const int head = kTail ? h0 + (g << pshift) + (row_in_tile & ((1 << pshift) - 1)) : h0 + g;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + lane_base + kColS + (uint32_t)(g * 128);
    const uint32_t tO = tmem_base + lane_base + kColO + (uint32_t)(g * 128);
    const float sc = p.scale_log2;                 // > 0 (checked by the launcher)
    float m_ref = -INFINITY, l_run = 0.f;          // running max (scaled, log2 domain) and row sum
    const uint32_t* bm_row = p.bitmask ? p.bitmask + (long long)min(i, n - 1) * p.bitmask_stride : nullptr;

    const uint64_t sc2 = pk_f2(sc, sc);

    for (int j = 0; j < total && kDbg != 1 && !(kDbg == 3 && g == 1); ++j) {
      if constexpr (kAhead) ptx::mbar_wait(bar(bPV0Done + g), j & 1);          // first 64 keys of S(j)
      else ptx::mbar_wait(bar(bSFull + g), kDbg == 2 ? 0 : (j & 1));
      ptx::tc_fence_after();
      const int c0 = tile_of(j) * kBN;
      const bool full = bm_row == nullptr &&
                        tile_is_full(r0, r1, c0, c0 + kBN - 1, n, lrow, p.si, p.sl, p.mode);
      if constexpr (kSplitP) {
        // Two halves of 64 keys, each an optimistic pass of its own against the current reference
        // max.  A refresh in the second half rescales O after PV's first half (already issued
        // with the old reference) has completed: O, l and the first half's sum all scale by alpha.
        float sum_a = 0.f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t ph[32];           // 64 probabilities as bf16 pairs
          float sum_h;
#pragma unroll 1
          for (int attempt = 0;; ++attempt) {
            const float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
            const uint64_t negm2 = pk_f2(neg_m, neg_m);
            float mx0 = -INFINITY, mx1 = -INFINITY;
            uint64_t sum2 = pk_f2(0.f, 0.f);
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
                ph[(c & 1) * 16 + e / 2] = pack2(a, b);
              }
            };
            uint32_t xa[32], xb[32];
            ptx::tmem_ld32(tS + h * 64, xa);
            ptx::tc_wait_ld();
            ptx::tmem_ld32(tS + h * 64 + 32, xb);      // in flight while the first chunk is processed
            chunk(2 * h, xa);
            ptx::tc_wait_ld();
            chunk(2 * h + 1, xb);
            const float m_new = fmaxf(m_ref, fmaxf(mx0, mx1) * sc);
            const bool grow = (m_new > m_ref + 8.0f) || (m_ref == -INFINITY && m_new > -INFINITY);
            {
              float a, b;
              upk_f2(sum2, a, b);
              sum_h = a + b;
            }
            if (attempt == 1 || !__any_sync(0xffffffffu, grow)) break;
            if (j > 0 || h == 1) {
              // h == 0: S(j) complete => PV(j-1) complete; h == 1: wait for the first half of PV(j)
              if (h == 1 && kDbg != 2) { ptx::mbar_wait(bar(bPV0Done + g), j & 1); ptx::tc_fence_after(); }
              const float alpha = (m_new == -INFINITY) ? 1.0f : fast_exp2(m_ref - m_new);
              l_run *= alpha;
              sum_a *= alpha;
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
          if (h == 0) sum_a = sum_h; else l_run += sum_a + sum_h;
          ptx::tmem_st32(tS + h * 32, ph);         // S columns [0,64) were consumed by the first half
          ptx::tc_wait_st();
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0 && kDbg != 2) ptx::mbar_arrive(bar((h == 0 ? bPReady : bP1Ready) + g));
        }
        continue;
      }
      if constexpr (kPipe) {
        uint32_t pk[64];
        uint64_t sum2;
        // column blocks of this tile's two 64-key halves (kAhead: they swap every tile); P goes to H0's
        const uint32_t tH0 = kAhead ? tS + (uint32_t)((j & 1) * 64) : tS;
        const uint32_t tH1 = kAhead ? tS + (uint32_t)(((j & 1) ^ 1) * 64) : tS + 64;
#pragma unroll 1
        for (int attempt = 0;; ++attempt) {
          const float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
          const uint64_t negm2 = pk_f2(neg_m, neg_m);
          float mx0 = -INFINITY, mx1 = -INFINITY;
          sum2 = pk_f2(0.f, 0.f);
          // A: raw scores of chunk c (in xr) -> t = x * scale - m_ref, in place; row max tracked
          auto phase_a = [&](int c, uint32_t (&xr)[32]) {
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
              float a, b;
              upk_f2(ffma2(pk_f2(x[e], x[e + 1]), sc2, negm2), a, b);
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
          uint32_t b0[32], b1[32], b2[32];
          ptx::tmem_ld32(tH0, b0);
          ptx::tc_wait_ld();
          ptx::tmem_ld32(tH0 + 32, b1);
          phase_a(0, b0);
          ptx::tc_wait_ld();
          if constexpr (kAhead) {
            // everything that only needs H0 first (half of the pass), then wait for keys 64..127
            phase_a(1, b1);
            phase_b(0, b0);
            phase_b(1, b1);
            ptx::mbar_wait(bar(bSFull + g), j & 1);   // (completed long ago on a redo)
            ptx::tc_fence_after();
            ptx::tmem_ld32(tH1, b2);
            ptx::tc_wait_ld();
            ptx::tmem_ld32(tH1 + 32, b0);
            phase_a(2, b2);
          } else {
            ptx::tmem_ld32(tH1, b2);
            phase_a(1, b1);
            phase_b(0, b0);
            ptx::tc_wait_ld();
            ptx::tmem_ld32(tH1 + 32, b0);
            phase_a(2, b2);
            phase_b(1, b1);
          }
          ptx::tc_wait_ld();
          phase_a(3, b0);
          phase_b(2, b2);
          const float m_new = fmaxf(m_ref, fmaxf(mx0, mx1) * sc);
          const bool grow = (m_new > m_ref + 8.0f) || (m_ref == -INFINITY && m_new > -INFINITY);
          if (attempt == 1 || !__any_sync(0xffffffffu, grow)) {
            if constexpr (kAhead) {                   // H1(j)'s block is read and will not be re-read
              ptx::tc_fence_before();
              __syncwarp();
              if (lane == 0) ptx::mbar_arrive(bar(bP1Ready + g));
            }
            // every score has been read and the pass stands: the first half of P can go out
            // while the last chunk is still being exponentiated
            ptx::tmem_st32(tH0, *reinterpret_cast<uint32_t(*)[32]>(&pk[0]));
            phase_b(3, b0);
            break;
          }
          if (j > 0) {
            const float alpha = (m_new == -INFINITY) ? 1.0f : fast_exp2(m_ref - m_new);
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
        ptx::tmem_st32(tH0 + 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[32]));
        ptx::tc_wait_st();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0 && kDbg != 2) ptx::mbar_arrive(bar(bPReady + g));
        continue;
      }
      uint32_t pk[64];             // the row's 128 probabilities as bf16 pairs
      float mx;                    // row max of the raw scores of this tile
      uint64_t sum2;               // two partial row sums
      // Optimistic single pass over S: exponentiate against the CURRENT reference max while
      // tracking the tile's max.  Only if some row's max outgrew the reference by more than 2^8
      // (always on the first tile, rarely afterwards) the reference is refreshed, O and l are
      // rescaled and the pass is redone -- S is still intact in TMEM because P is stored last.
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
        mx = fmaxf(mx0, mx1);
        const float m_new = fmaxf(m_ref, mx * sc);
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
      if (lane == 0 && kDbg != 2) ptx::mbar_arrive(bar(bPReady + g));
    }
    // ---- epilogue: O / l -> bf16 -> global ------------------------------------
    if (!(kDbg == 3 && g == 1)) ptx::mbar_wait(bar(bOFull + g), 0);
    ptx::tc_fence_after();
    if constexpr (kTail) {
      // fp32 partial of this KV chunk: O (unnormalised), reference max (log2 domain), l
      const int tt = kBM >> pshift;
      const long long prow = (part * p.Hq + head) * tt + (i - r0);
      float* po = p.part_o + prow * kD;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t r[32];
        ptx::tmem_ld32(tO + c * 32, r);
        ptx::tc_wait_ld();
        if (i <= r1) {
#pragma unroll
          for (int v4 = 0; v4 < 8; ++v4)
            *reinterpret_cast<uint4*>(po + c * 32 + v4 * 4) =
                make_uint4(r[v4 * 4 + 0], r[v4 * 4 + 1], r[v4 * 4 + 2], r[v4 * 4 + 3]);
        }
      }
      if (i <= r1) p.part_ml[prow] = make_float2(m_ref, l_run);
    } else {
    const float inv = l_run > 0.f ? fast_rcp(l_run) : 0.f;
    __nv_bfloat16* orow = p.out + (long long)(seq_start + i) * p.o_stride + head * kD;
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
    }
    ptx::tc_fence_before();
  }

  // ---- teardown -------------------------------------------------------------------
  __syncthreads();
  if (warp == 9) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

// Merge the split-KV partials of the tail tiles: one warp per (row, head), 4 columns per lane.
__global__ void __launch_bounds__(256) combine_tail_kernel(const AttnParams p, const int tt) {
  constexpr int kRowsPerCta = 8;
  const int rblks = tt / kRowsPerCta;
  int e = blockIdx.x / rblks;
  const int rblk = blockIdx.x % rblks;
  const int ql = e % p.last_tiles, seq = e / p.last_tiles;
  const int head = blockIdx.y;
  const int start = p.cu[seq], n = p.cu[seq + 1] - start;
  const int lrow = first_dense_row(n, p.last);
  if (n <= 0 || lrow >= n) return;
  const int r0 = lrow + ql * tt;
  if (r0 >= n) return;
  const int r1 = min(n, r0 + tt) - 1;
  const int ns = tail_nsplit(r1 / kBN + 1, p.nsplit_max);
  const long long base = (long long)(seq * p.last_tiles + ql) * p.nsplit_max;
  const int r = rblk * kRowsPerCta + (threadIdx.x >> 5), c4 = (threadIdx.x & 31) * 4;
  if (r0 + r > r1) return;
  float mmax = -INFINITY;
  for (int s = 0; s < ns; ++s)
    mmax = fmaxf(mmax, p.part_ml[((base + s) * p.Hq + head) * tt + r].x);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  float l = 0.f;
#pragma unroll 4
  for (int s = 0; s < ns; ++s) {
    const long long prow = ((base + s) * p.Hq + head) * tt + r;
    const float2 ml = p.part_ml[prow];
    const float w = (ml.x == -INFINITY) ? 0.f : fast_exp2(ml.x - mmax);
    l += ml.y * w;
    const float4 v = *reinterpret_cast<const float4*>(p.part_o + prow * kD + c4);
    acc.x += v.x * w; acc.y += v.y * w; acc.z += v.z * w; acc.w += v.w * w;
  }
  const float inv = l > 0.f ? 1.f / l : 0.f;
  __nv_bfloat16* og = p.out + (long long)(start + r0 + r) * p.o_stride + head * kD + c4;
  *reinterpret_cast<uint2*>(og) = make_uint2(pack2(acc.x * inv, acc.y * inv), pack2(acc.z * inv, acc.w * inv));
  if (p.lse != nullptr && c4 == 0)
    p.lse[(long long)head * p.total_T + start + r0 + r] = mmax * 0.6931471805599453f + logf(l);
}

}  // namespace

// Query-tile size (tokens) of the tcgen05 tail path: 2 groups x 128 packed rows = all G heads.
int tail_tc_block_m(int Hq, int Hkv) { return 256 / (Hq / Hkv); }
bool tail_tc_supported(int Hq, int Hkv, float scale_log2) {
  if (Hkv <= 0 || Hq % Hkv != 0) return false;
  const int G = Hq / Hkv;
  return (G == 2 || G == 4 || G == 8) && scale_log2 > 0.f;
}

// K/V descriptors: a contiguous [rows, Hkv * D] buffer (128-row boxes) or the paged cache viewed as
// [num_blocks * page_size, Hkv * D] (one box per page when pages are shorter than a tile).
static int kv_maps(TmaMaps* maps, const AttnParams& p) {
  const int box = (p.block_table != nullptr && p.page_size < kBN) ? p.page_size : kBN;
  int st = tma_map_2d(&maps->k, p.k, p.total_kv, (long long)p.Hkv * kD, p.kv_stride, box);
  if (st == TMX_OK) st = tma_map_2d(&maps->v, p.v, p.total_kv, (long long)p.Hkv * kD, p.kv_stride, box);
  return st;
}

// Dense tail rows of a triangle layer: split-KV workers on the tensor cores + merge.
// p.tail_bm = tail_tc_block_m(), p.last_tiles / p.nsplit_max from the plan.
int launch_attn_tail_tc(const AttnParams& p, cudaStream_t s) {
  const int G = p.Hq / p.Hkv, P = G / 2, TT = kBM / P;
  auto kern = attn_sm100_v3_kernel<4, 0, true, 0>;
  TMX_SET_SMEM_ATTR(kern, kAttrV3Tail, kSmem);
  TmaMaps maps;
  int st = tma_map_q3d(&maps.q, p.q, p.total_T, p.Hq, p.q_stride, P, TT);
  if (st == TMX_OK) st = kv_maps(&maps, p);
  if (st != TMX_OK) return st;
  dim3 grid((unsigned)(p.B * p.last_tiles * p.nsplit_max), (unsigned)p.Hkv);
  kern<<<grid, kThreads, kSmem, s>>>(p, maps);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  dim3 cg((unsigned)(p.B * p.last_tiles * (TT / 8)), (unsigned)p.Hq);
  combine_tail_kernel<<<cg, 256, 0, s>>>(p, TT);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

int launch_attn_sm100_v3(const AttnParams& p, int variant, cudaStream_t s) {
  if ((p.Hq / p.Hkv) % 2 != 0) return TMX_ERR_UNSUPPORTED;   // needs two query heads per KV head
  if (!(p.scale_log2 > 0.f)) return TMX_ERR_UNSUPPORTED;     // row max is taken on raw scores
  using Kern = void (*)(const AttnParams, const TmaMaps);
  // Product kernel: the look-ahead variant with every exponential on MUFU (+1.2 % over the classic
  // single-MMA-per-tile kernel at 128K, profiles/r01_sm100_v3_experiments.md section 7).
  Kern kern = attn_sm100_v3_kernel<0, 0, false, 3>;
  int slot = kAttrV3Dense;
#ifdef TMX_DEV
  // Developer build only (make DEV=1): timing / emulation variants selected by TMX_SM100_VARIANT.
  // bits 0-1: exponentials on the FMA pipe per 16 (0 = 4, 1 = none, 2 = 6, 3 = 8); bits 2-3: kDbg
  // (results are garbage); bit 4: kSplitP; bit 6: pipelined pass; 6+4: look-ahead; 6+5+4: polling issuer.
  if (variant != 0) {
    const int emu = variant & 3, dbg = (variant >> 2) & 3;
    static const Kern table[4][4] = {
        {attn_sm100_v3_kernel<4, 0, false, 0>, attn_sm100_v3_kernel<0, 0, false, 0>, attn_sm100_v3_kernel<6, 0, false, 0>, attn_sm100_v3_kernel<8, 0, false, 0>},
        {attn_sm100_v3_kernel<0, 1, false, 0>, attn_sm100_v3_kernel<0, 1, false, 0>, attn_sm100_v3_kernel<0, 1, false, 0>, attn_sm100_v3_kernel<0, 1, false, 0>},
        {attn_sm100_v3_kernel<0, 2, false, 0>, attn_sm100_v3_kernel<4, 2, false, 0>, attn_sm100_v3_kernel<6, 2, false, 0>, attn_sm100_v3_kernel<8, 2, false, 0>},
        {attn_sm100_v3_kernel<0, 3, false, 0>, attn_sm100_v3_kernel<4, 3, false, 0>, attn_sm100_v3_kernel<6, 3, false, 0>, attn_sm100_v3_kernel<8, 3, false, 0>}};
    kern = table[dbg][emu];
    if (dbg == 0 && (variant & 16)) {
      static const Kern split[4] = {attn_sm100_v3_kernel<4, 0, false, 1>, attn_sm100_v3_kernel<0, 0, false, 1>,
                                    attn_sm100_v3_kernel<6, 0, false, 1>, attn_sm100_v3_kernel<8, 0, false, 1>};
      kern = split[emu];
    }
    if (dbg == 0 && (variant & 64)) {
      static const Kern pipe[4] = {attn_sm100_v3_kernel<4, 0, false, 2>, attn_sm100_v3_kernel<0, 0, false, 2>,
                                   attn_sm100_v3_kernel<6, 0, false, 2>, attn_sm100_v3_kernel<8, 0, false, 2>};
      static const Kern ahead[4] = {attn_sm100_v3_kernel<4, 0, false, 3>, attn_sm100_v3_kernel<0, 0, false, 3>,
                                    attn_sm100_v3_kernel<6, 0, false, 3>, attn_sm100_v3_kernel<8, 0, false, 3>};
      static const Kern poll[4] = {attn_sm100_v3_kernel<4, 0, false, 4>, attn_sm100_v3_kernel<0, 0, false, 4>,
                                   attn_sm100_v3_kernel<6, 0, false, 4>, attn_sm100_v3_kernel<8, 0, false, 4>};
      kern = (variant & 16) ? ((variant & 32) ? poll[emu] : ahead[emu]) : pipe[emu];
    }
    slot = -1;
  }
#else
  (void)variant;
#endif
  if (slot < 0) TMX_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
  else TMX_SET_SMEM_ATTR(kern, slot, kSmem);
  TmaMaps maps;
  int st = tma_map_2d(&maps.q, p.q, p.total_T, (long long)p.Hq * kD, p.q_stride, kBM);
  if (st == TMX_OK) st = kv_maps(&maps, p);
  if (st != TMX_OK) return st;
  dim3 grid((unsigned)p.n_slots, (unsigned)(p.Hq / 2));
  kern<<<grid, kThreads, kSmem, s>>>(p, maps);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

}  // namespace tmx

// Device code of the tcgen05 attention kernel (included by tmx_attn_sm100.cu).
#pragma once

namespace tmx {
namespace {

// 2^t for t <= ~8 on the FMA/ALU pipes (no MUFU): round-to-nearest split t = n + f,
// f in [-0.5, 0.5], cubic minimax for 2^f (max rel. error 7.5e-5, far below bf16's 2^-9),
// then n is added into the exponent field.
__device__ __forceinline__ float exp2_poly(float t) {
  t = fmaxf(t, -126.0f);
  const float r = t + 12582912.0f;          // 1.5 * 2^23: low mantissa bits of r hold n
  const float f = t - (r - 12582912.0f);
  float pl = fmaf(0.0551716685f, f, 0.2426111251f);
  pl = fmaf(pl, f, 0.6932609677f);
  pl = fmaf(pl, f, 0.9999280572f);
  return __int_as_float(__float_as_int(pl) + (__float_as_int(r) << 23));
}

// kDbg (developer timing experiments, results are garbage): 1 = MMA/TMA side alone (no softmax,
// no waits for P), 2 = softmax side alone (S tiles 0/1 computed once and re-read).
// kDbg 3 = as 1, but S issued as N=128 MMAs every other tile (timing of the MMA shape only).
// kPing: the two softmax warps that share an SM sub-partition take turns in their exp phase.
template <int kEmu, int kDbg = 0, bool kPing = false>
__global__ void __launch_bounds__(kThreads, 1)
attn_sm100_kernel(const AttnParams p, const __grid_constant__ TmaMaps maps) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;   // SWIZZLE_128B: 1024-byte atoms
  const uint32_t sQ = smem_base, sK = smem_base + kOffK, sV = smem_base + kOffV;
  const uint32_t sBar = smem_base + kOffBar;
  const uint32_t sTmemPtr = sBar + kNumBars * 8;
  auto bar = [&](int i) { return sBar + 8u * (uint32_t)i; };

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // provably warp-uniform
  const int lane = threadIdx.x & 31;

  // ---- work item (uniform per CTA) -------------------------------------------
  const SlotInfo slot = p.slots[p.n_slots - 1 - blockIdx.x];      // descending: longest first
  if (slot.seq < 0) return;
  const int n = slot.n, r0 = slot.q_tile * kBM;
  const int r1 = min(n, r0 + kBM) - 1;
  const int lo0 = slot.lo0, cnt0 = slot.hi0 - slot.lo0, lo1 = slot.lo1;
  const int total = cnt0 + (slot.hi1 - slot.lo1);
  const int seq_start = p.cu[slot.seq];
  const int h0 = 2 * blockIdx.y;
  const int kvh = h0 / (p.Hq / p.Hkv);
  const int lrow = first_dense_row(n, p.last);
  auto tile_of = [&](int it) { return it < cnt0 ? lo0 + it : lo1 + (it - cnt0); };

  // ---- setup -------------------------------------------------------------------
  if (warp == 8 && lane == 0) {
    ptx::prefetch_tmap(&maps.q); ptx::prefetch_tmap(&maps.k); ptx::prefetch_tmap(&maps.v);
    ptx::mbar_init(bar(bQFull), 1);
    for (int s = 0; s < kKStages; ++s) { ptx::mbar_init(bar(bKFull + s), 1); ptx::mbar_init(bar(bKEmpty + s), 1); }
    for (int s = 0; s < kVStages; ++s) { ptx::mbar_init(bar(bVFull + s), 1); ptx::mbar_init(bar(bVEmpty + s), 1); }
    for (int i = 0; i < 4; ++i) {
      ptx::mbar_init(bar(bSFull + i), 1);
      ptx::mbar_init(bar(bPReady + i), 4);       // one arrive per softmax warp of the group
    }
    for (int g = 0; g < 2; ++g) { ptx::mbar_init(bar(bPVDone + g), 1); ptx::mbar_init(bar(bOFull + g), 1); }
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
    // Load order follows consumption: K(0) K(1) { V(j) K(j+2) }_j
    if (ptx::elect_one()) {
      ptx::mbar_expect_tx(bar(bQFull), kQBytes);
      for (int g = 0; g < 2; ++g)
        for (int pnl = 0; pnl < 2; ++pnl)
          ptx::tma_load_2d(sQ + (g * 2 + pnl) * (kBM * 128), &maps.q, bar(bQFull),
                           (h0 + g) * kD + pnl * 64, seq_start + r0);
      auto load_k = [&](int it) {
        const int st = it % kKStages, round = it / kKStages;
        if (round > 0) ptx::mbar_wait(bar(bKEmpty + st), (round - 1) & 1);
        ptx::mbar_expect_tx(bar(bKFull + st), kTileBytes);
        const int row = seq_start + tile_of(it) * kBN;
        for (int pnl = 0; pnl < 2; ++pnl)
          ptx::tma_load_2d(sK + st * kTileBytes + pnl * kPanelBytes, &maps.k, bar(bKFull + st),
                           kvh * kD + pnl * 64, row);
      };
      auto load_v = [&](int it) {
        const int st = it % kVStages, round = it / kVStages;
        if (round > 0) ptx::mbar_wait(bar(bVEmpty + st), (round - 1) & 1);
        ptx::mbar_expect_tx(bar(bVFull + st), kTileBytes);
        const int row = seq_start + tile_of(it) * kBN;
        for (int pnl = 0; pnl < 2; ++pnl)
          ptx::tma_load_2d(sV + st * kTileBytes + pnl * kPanelBytes, &maps.v, bar(bVFull + st),
                           kvh * kD + pnl * 64, row);
      };
      load_k(0);
      if (total > 1) load_k(1);
      for (int it = 0; it < total && kDbg != 2; ++it) {
        load_v(it);
        if (it + 2 < total) load_k(it + 2);
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // =========================== MMA issuer ===========================
    // The whole warp runs the loop (warp-uniform control flow keeps descriptors and
    // addresses in uniform registers); one elected lane issues the tcgen05 instructions.
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const int ntot = __shfl_sync(0xffffffffu, total, 0);
    constexpr uint32_t kIdescS = ptx::instr_desc_bf16(128, kBN, false);
    constexpr uint32_t kIdescO = ptx::instr_desc_bf16(128, 128, true);
    const uint64_t dQ = ptx::smem_desc(sQ, 16, 1024);
    const uint64_t dK = ptx::smem_desc(sK, 16, 1024);
    const uint64_t dV = ptx::smem_desc(sV, kPanelBytes, 1024);

    // S_g(tile in K stage kst) -> S buffer `buf`: 8 k-steps of 16 over d (panels of 64 d)
    auto issue_s = [&](int g, int buf, int kst) {
      const uint32_t d_tmem = tb + kColS + (uint32_t)(g * 128 + buf * 64);
      const uint64_t qa = dQ + (uint64_t)((g * 2 * (kBM * 128)) >> 4);
      const uint64_t kb = dK + (uint64_t)((kst * kTileBytes) >> 4);
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {
        const uint32_t qo = ((ks >> 2) * (kBM * 128) + (ks & 3) * 32) >> 4;
        const uint32_t ko = ((ks >> 2) * kPanelBytes + (ks & 3) * 32) >> 4;
        ptx::mma_ss(d_tmem, qa + qo, kb + ko, kIdescS, ks > 0);
      }
    };
    // O_g += P_g V: 4 k-steps of 16 keys; P (bf16) read from the S buffer it overwrote
    auto issue_pv = [&](int g, int buf, int vst, bool acc) {
      const uint32_t d_tmem = tb + kColO + (uint32_t)(g * 128);
      const uint32_t a_tmem = tb + kColS + (uint32_t)(g * 128 + buf * 64);
      const uint64_t vb = dV + (uint64_t)((vst * kTileBytes) >> 4);
#pragma unroll
      for (int ks = 0; ks < kBN / 16; ++ks)
        ptx::mma_ts(d_tmem, a_tmem + (uint32_t)(ks * 8), vb + (uint64_t)((ks * 2048) >> 4), kIdescO,
                    acc || ks > 0);
    };

    ptx::mbar_wait(bar(bQFull), 0);
    for (int t = 0; t < 2 && t < ntot; ++t) {
      ptx::mbar_wait(bar(bKFull + t), 0);
      ptx::tc_fence_after();
      if (ptx::elect_one()) {
        issue_s(0, t, t); ptx::mma_commit(bar(bSFull + 0 + t));
        issue_s(1, t, t); ptx::mma_commit(bar(bSFull + 2 + t));
        ptx::mma_commit(bar(bKEmpty + t));
      }
      __syncwarp();
    }
    int vst = 0, vph = 0, kst = 2 % kKStages, kph = 0;     // stage / parity of V(j) and K(j+2)
    if (kDbg == 2 && ptx::elect_one()) { ptx::mma_commit(bar(bOFull + 0)); ptx::mma_commit(bar(bOFull + 1)); }
    constexpr uint32_t kIdescS128 = ptx::instr_desc_bf16(128, 128, false);
    for (int j = 0; j < ntot && kDbg != 2; ++j) {
      const int buf = j & 1, pph = (j >> 1) & 1;
      const bool more = j + 2 < ntot;
      ptx::mbar_wait(bar(bVFull + vst), vph);
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        if (kDbg != 1 && kDbg != 3) ptx::mbar_wait(bar(bPReady + g * 2 + buf), pph);
        if (more && g == 0) ptx::mbar_wait(bar(bKFull + kst), kph);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
          issue_pv(g, buf, vst, j > 0);
          ptx::mma_commit(bar(j + 1 < ntot ? bPVDone + g : bOFull + g));
          if (g == 1) ptx::mma_commit(bar(bVEmpty + vst));
          if (more && kDbg == 3) {
            if (buf == 0) {
              const uint32_t d_tmem = tb + kColS + (uint32_t)(g * 128);
              const uint64_t qa = dQ + (uint64_t)((g * 2 * (kBM * 128)) >> 4);
#pragma unroll
              for (int ks = 0; ks < 8; ++ks)
                ptx::mma_ss(d_tmem, qa + (uint64_t)(((ks >> 2) * (kBM * 128) + (ks & 3) * 32) >> 4),
                            dK + (uint64_t)(((ks >> 2) * kPanelBytes + (ks & 3) * 32) >> 4), kIdescS128, ks > 0);
            }
            ptx::mma_commit(bar(bSFull + g * 2 + buf));
            if (g == 1) ptx::mma_commit(bar(bKEmpty + kst));
          } else if (more) {
            issue_s(g, buf, kst);
            ptx::mma_commit(bar(bSFull + g * 2 + buf));
            if (g == 1) ptx::mma_commit(bar(bKEmpty + kst));
          }
        }
        __syncwarp();
      }
      if (++vst == kVStages) { vst = 0; vph ^= 1; }
      if (++kst == kKStages) { kst = 0; kph ^= 1; }
    }
  } else {
    // =========================== softmax groups ===========================
    const int g = warp >> 2;                       // 0: head h0, 1: head h0 + 1
    const int row_in_tile = (warp & 3) * 32 + lane;
    const int i = r0 + row_in_tile;                // query row inside the sequence
    const int head = h0 + g;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tSg = tmem_base + lane_base + kColS + (uint32_t)(g * 128);
    const uint32_t tO = tmem_base + lane_base + kColO + (uint32_t)(g * 128);
    const float sc = p.scale_log2;                 // > 0 (checked by the launcher)
    float m_ref = -INFINITY, l_run = 0.f;          // running max (scaled, log2 domain) and row sum
    const uint32_t* bm_row = p.bitmask ? p.bitmask + (long long)min(i, n - 1) * p.bitmask_stride : nullptr;

    const int tok_mine = 1 + (warp & 3) * 2 + g, tok_other = 1 + (warp & 3) * 2 + (g ^ 1);
    if (kPing && g == 1) ptx::named_arrive(tok_other, 64);      // group A goes first
    for (int j = 0; j < total && kDbg != 1 && kDbg != 3; ++j) {
      const int buf = j & 1;
      const uint32_t tS = tSg + (uint32_t)(buf * 64);
      ptx::mbar_wait(bar(bSFull + g * 2 + buf), kDbg == 2 ? 0 : (j >> 1) & 1);
      ptx::tc_fence_after();
      uint32_t xr[kBN];
      ptx::tmem_ld32(tS, *reinterpret_cast<uint32_t(*)[32]>(&xr[0]));
      ptx::tmem_ld32(tS + 32, *reinterpret_cast<uint32_t(*)[32]>(&xr[32]));
      ptx::tc_wait_ld();
      float x[kBN];
#pragma unroll
      for (int e = 0; e < kBN; ++e) x[e] = __uint_as_float(xr[e]);

      const int c0 = tile_of(j) * kBN;
      const bool full = bm_row == nullptr &&
                        tile_is_full(r0, r1, c0, c0 + kBN - 1, n, lrow, p.si, p.sl, p.mode);
      if (!full) {
        if (bm_row == nullptr) {
#pragma unroll
          for (int e = 0; e < kBN; ++e)
            x[e] = attends(i, c0 + e, n, lrow, p.si, p.sl, p.mode) ? x[e] : -INFINITY;
        } else {
#pragma unroll
          for (int w = 0; w < kBN / 32; ++w) {
            const uint32_t bits = bm_row[(c0 >> 5) + w];
#pragma unroll
            for (int e = 0; e < 32; ++e) x[w * 32 + e] = ((bits >> e) & 1u) ? x[w * 32 + e] : -INFINITY;
          }
        }
      }
      float mx0 = x[0], mx1 = x[1], mx2 = x[2], mx3 = x[3];
#pragma unroll
      for (int e = 4; e < kBN; e += 4) {
        mx0 = fmaxf(mx0, x[e]); mx1 = fmaxf(mx1, x[e + 1]);
        mx2 = fmaxf(mx2, x[e + 2]); mx3 = fmaxf(mx3, x[e + 3]);
      }
      const float m_new = fmaxf(m_ref, fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * sc);
      // refresh the reference max lazily: only when some row of the warp grew by > 2^8
      const bool grow = (m_new > m_ref + 8.0f) || (m_ref == -INFINITY && m_new > -INFINITY);
      if (__any_sync(0xffffffffu, grow)) {
        if (j > 0) {
          // O holds tiles 0..j-1 once PV(j-1) has completed; PV(j) cannot start before our arrive
          if (kDbg != 2) ptx::mbar_wait(bar(bPVDone + g), (j - 1) & 1);
          ptx::tc_fence_after();
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
      const float neg_m = (m_ref == -INFINITY) ? 0.f : -m_ref;
      if (kPing) ptx::named_sync(tok_mine, 64);
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
      uint32_t pk[kBN / 2];
#pragma unroll
      for (int e = 0; e < kBN; e += 4) {
        float q0 = fmaf(x[e], sc, neg_m), q1 = fmaf(x[e + 1], sc, neg_m);
        float q2 = fmaf(x[e + 2], sc, neg_m), q3 = fmaf(x[e + 3], sc, neg_m);
        q0 = ((e & 7) + 0 < kEmu) ? exp2_poly(q0) : fast_exp2(q0);
        q1 = ((e & 7) + 1 < kEmu) ? exp2_poly(q1) : fast_exp2(q1);
        q2 = ((e & 7) + 2 < kEmu) ? exp2_poly(q2) : fast_exp2(q2);
        q3 = ((e & 7) + 3 < kEmu) ? exp2_poly(q3) : fast_exp2(q3);
        s0 += q0; s1 += q1; s2 += q2; s3 += q3;
        pk[e / 2] = pack2(q0, q1);
        pk[e / 2 + 1] = pack2(q2, q3);
      }
      l_run += (s0 + s1) + (s2 + s3);
      if (kPing) ptx::named_arrive(tok_other, 64);
      ptx::tmem_st32(tS, pk);
      ptx::tc_wait_st();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0 && kDbg != 2) ptx::mbar_arrive(bar(bPReady + g * 2 + buf));
    }

    // ---- epilogue: O / l -> bf16 -> global ------------------------------------
    // (a parity wait on bPVDone could pass one completion early here: PV(total-2) may still be in flight)
    ptx::mbar_wait(bar(bOFull + g), 0);
    ptx::tc_fence_after();
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
    ptx::tc_fence_before();
  }

  // ---- teardown -------------------------------------------------------------------
  __syncthreads();
  if (warp == 9) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace
}  // namespace tmx

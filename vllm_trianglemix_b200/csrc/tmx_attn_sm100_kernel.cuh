// Device code of the tcgen05 attention kernel (included by tmx_attn_sm100.cu).
#pragma once

namespace tmx {
namespace {

template <int BN, bool kPTmem>
__global__ void __launch_bounds__(kThreads, 1)
attn_sm100_kernel(const AttnParams p, const __grid_constant__ TmaMaps maps) {
  using C = Cfg<BN, kPTmem>;
  using B = Bars<BN, kPTmem>;
  constexpr int S = C::kStages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;   // SWIZZLE_128B: 1024-byte atoms
  const uint32_t sQ = smem_base, sK = smem_base + C::kOffK, sV = smem_base + C::kOffV;
  const uint32_t sP = smem_base + C::kOffP, sBar = smem_base + C::kOffBar;
  const uint32_t sTmemPtr = sBar + C::kNumBars * 8;
  auto bar = [&](int i) { return sBar + 8u * (uint32_t)i; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

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
    ptx::mbar_init(bar(B::q_full), 1);
    for (int s = 0; s < S; ++s) {
      ptx::mbar_init(bar(B::k_full + s), 1); ptx::mbar_init(bar(B::v_full + s), 1);
      ptx::mbar_init(bar(B::k_empty + s), 1); ptx::mbar_init(bar(B::v_empty + s), 1);
    }
    for (int g = 0; g < 2; ++g) {
      ptx::mbar_init(bar(B::s_full + g), 1);
      ptx::mbar_init(bar(B::p_ready + g), 128);
      ptx::mbar_init(bar(B::o_full + g), 1);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 9) ptx::tmem_alloc(sTmemPtr, 512);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(sTmemPtr));

  constexpr uint32_t kIdescS = ptx::instr_desc_bf16(128, BN, false);
  constexpr uint32_t kIdescO = ptx::instr_desc_bf16(128, 128, true);

  if (warp == 8) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      ptx::mbar_expect_tx(bar(B::q_full), kQBytes);
      for (int g = 0; g < 2; ++g)
        for (int pnl = 0; pnl < 2; ++pnl)
          ptx::tma_load_2d(sQ + (g * 2 + pnl) * (kBM * 128), &maps.q, bar(B::q_full),
                           (h0 + g) * kD + pnl * 64, seq_start + r0);
      for (int it = 0; it < total; ++it) {
        const int st = it % S, round = it / S;
        const int row = seq_start + tile_of(it) * BN;
        if (round > 0) ptx::mbar_wait(bar(B::k_empty + st), (round - 1) & 1);
        ptx::mbar_expect_tx(bar(B::k_full + st), C::kTileBytes);
        for (int pnl = 0; pnl < 2; ++pnl)
          ptx::tma_load_2d(sK + st * C::kTileBytes + pnl * C::kPanelBytes, &maps.k, bar(B::k_full + st),
                           kvh * kD + pnl * 64, row);
        if (round > 0) ptx::mbar_wait(bar(B::v_empty + st), (round - 1) & 1);
        ptx::mbar_expect_tx(bar(B::v_full + st), C::kTileBytes);
        for (int pnl = 0; pnl < 2; ++pnl)
          ptx::tma_load_2d(sV + st * C::kTileBytes + pnl * C::kPanelBytes, &maps.v, bar(B::v_full + st),
                           kvh * kD + pnl * 64, row);
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // =========================== MMA issuer ===========================
    if (lane == 0) {
      // S_g(j) = Q_g K(j)^T : 8 k-steps of 16 over d; Q/K panels hold 64 d each (4 k-steps per panel)
      auto issue_s = [&](int g, int st) {
        const uint32_t d_tmem = tmem_base + (uint32_t)(g * 128);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint32_t qa = sQ + (g * 2 + (ks >> 2)) * (kBM * 128) + (ks & 3) * 32;
          const uint32_t kb = sK + st * C::kTileBytes + (ks >> 2) * C::kPanelBytes + (ks & 3) * 32;
          ptx::mma_ss(d_tmem, ptx::smem_desc(qa, 16, 1024), ptx::smem_desc(kb, 16, 1024), kIdescS, ks > 0);
        }
      };
      // O_g += P_g V(j) : BN/16 k-steps over keys; V is MN-major (d contiguous): panels of 64 d
      auto issue_pv = [&](int g, int st, bool acc) {
        const uint32_t d_tmem = tmem_base + 256u + (uint32_t)(g * 128);
#pragma unroll
        for (int ks = 0; ks < BN / 16; ++ks) {
          const uint32_t vb = sV + st * C::kTileBytes + ks * 2048;
          const uint64_t bdesc = ptx::smem_desc(vb, C::kPanelBytes, 1024);
          if (kPTmem) {
            ptx::mma_ts(d_tmem, tmem_base + (uint32_t)(g * 128 + ks * 8), bdesc, kIdescO, acc || ks > 0);
          } else {
            const uint32_t pa = sP + g * (C::kPBytes / 2) + (ks >> 2) * (kBM * 128) + (ks & 3) * 32;
            ptx::mma_ss(d_tmem, ptx::smem_desc(pa, 16, 1024), bdesc, kIdescO, acc || ks > 0);
          }
        }
      };
      ptx::mbar_wait(bar(B::q_full), 0);
      ptx::mbar_wait(bar(B::k_full + 0), 0);
      ptx::tc_fence_after();
      issue_s(0, 0); ptx::mma_commit(bar(B::s_full + 0));
      issue_s(1, 0); ptx::mma_commit(bar(B::s_full + 1));
      ptx::mma_commit(bar(B::k_empty + 0));
      for (int j = 0; j < total; ++j) {
        const int st = j % S, ph = (j / S) & 1;
        const int st1 = (j + 1) % S, ph1 = ((j + 1) / S) & 1;
        const bool more = j + 1 < total;
        ptx::mbar_wait(bar(B::v_full + st), ph);
        for (int g = 0; g < 2; ++g) {
          ptx::mbar_wait(bar(B::p_ready + g), j & 1);
          ptx::tc_fence_after();
          issue_pv(g, st, j > 0);
          if (g == 1) ptx::mma_commit(bar(B::v_empty + st));
          if (more) {
            if (g == 0) { ptx::mbar_wait(bar(B::k_full + st1), ph1); ptx::tc_fence_after(); }
            issue_s(g, st1);
            ptx::mma_commit(bar(B::s_full + g));
            if (g == 1) ptx::mma_commit(bar(B::k_empty + st1));
          } else {
            ptx::mma_commit(bar(B::o_full + g));
          }
        }
      }
    }
    __syncwarp();
  } else {
    // =========================== softmax groups ===========================
    const int g = warp >> 2;                       // 0: head h0, 1: head h0 + 1
    const int row_in_tile = (warp & 3) * 32 + lane;
    const int i = r0 + row_in_tile;                // query row inside the sequence
    const int head = h0 + g;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tS = tmem_base + lane_base + (uint32_t)(g * 128);
    const uint32_t tO = tmem_base + lane_base + 256u + (uint32_t)(g * 128);
    const float sc = p.scale_log2;
    float m_ref = -INFINITY, l_run = 0.f;
    const uint32_t* bm_row = p.bitmask ? p.bitmask + (long long)min(i, n - 1) * p.bitmask_stride : nullptr;

    for (int j = 0; j < total; ++j) {
      ptx::mbar_wait(bar(B::s_full + g), j & 1);
      ptx::tc_fence_after();
      const int c0 = tile_of(j) * BN;
      const bool full = bm_row == nullptr &&
                        tile_is_full(r0, r1, c0, c0 + BN - 1, n, lrow, p.si, p.sl, p.mode);
      float x[BN];
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        ptx::tmem_ld32(tS + c * 32, r);
        ptx::tc_wait_ld();
#pragma unroll
        for (int e = 0; e < 32; ++e) x[c * 32 + e] = __uint_as_float(r[e]) * sc;
      }
      if (!full) {
        if (bm_row == nullptr) {
#pragma unroll
          for (int e = 0; e < BN; ++e)
            x[e] = attends(i, c0 + e, n, lrow, p.si, p.sl, p.mode) ? x[e] : -INFINITY;
        } else {
#pragma unroll
          for (int w = 0; w < BN / 32; ++w) {
            const uint32_t bits = bm_row[(c0 >> 5) + w];
#pragma unroll
            for (int e = 0; e < 32; ++e) x[w * 32 + e] = ((bits >> e) & 1u) ? x[w * 32 + e] : -INFINITY;
          }
        }
      }
      float mx = -INFINITY;
#pragma unroll
      for (int e = 0; e < BN; ++e) mx = fmaxf(mx, x[e]);
      const float m_new = fmaxf(m_ref, mx);
      // refresh the reference max lazily: only when some row of the warp grew by > 2^8
      const bool grow = (m_new > m_ref + 8.0f) || (m_ref == -INFINITY && m_new > -INFINITY);
      if (__any_sync(0xffffffffu, grow)) {
        if (j > 0) {
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
      const float m_use = (m_ref == -INFINITY) ? 0.f : m_ref;
      float sum = 0.f;
#pragma unroll
      for (int e = 0; e < BN; ++e) {
        x[e] = fast_exp2(x[e] - m_use);
        sum += x[e];
      }
      l_run += sum;
      if (kPTmem) {
#pragma unroll
        for (int c = 0; c < BN / 64; ++c) {
          uint32_t r[32];
#pragma unroll
          for (int e = 0; e < 32; ++e) r[e] = pack2(x[c * 64 + 2 * e], x[c * 64 + 2 * e + 1]);
          ptx::tmem_st32(tS + c * 32, r);
        }
        ptx::tc_wait_st();
      } else {
        // K-major, 128B swizzle: row r of a 64-key panel = 128 B; 16-byte chunk c lands at c ^ (r & 7)
        const uint32_t prow = sP + g * (C::kPBytes / 2) + (row_in_tile >> 3) * 1024 + (row_in_tile & 7) * 128;
#pragma unroll
        for (int c = 0; c < BN / 8; ++c) {
          const uint32_t addr = prow + (c >> 3) * (kBM * 128) + (((c & 7) ^ (row_in_tile & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr),
                       "r"(pack2(x[c * 8 + 0], x[c * 8 + 1])), "r"(pack2(x[c * 8 + 2], x[c * 8 + 3])),
                       "r"(pack2(x[c * 8 + 4], x[c * 8 + 5])), "r"(pack2(x[c * 8 + 6], x[c * 8 + 7])) : "memory");
        }
        ptx::fence_proxy_async();
      }
      ptx::tc_fence_before();
      ptx::mbar_arrive(bar(B::p_ready + g));
    }

    // ---- epilogue: O / l -> bf16 -> global ------------------------------------
    ptx::mbar_wait(bar(B::o_full + g), 0);
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

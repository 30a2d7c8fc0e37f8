// Masked GQA prefill attention, warp-level tensor-core kernel (mma.sync m16n8k16
// bf16, fp32 accumulate).  Flash-style: one CTA per (query tile, query head),
// iterating ONLY the KV tiles the block index lists; online softmax in registers;
// P stays in registers between the two contractions.
//
// Role in the build: (1) the kernel for triangle layers at the reference's default
// parameters, where the work per query tile is 3-4 small KV tiles and the layer is
// HBM-bound (64 x 64 tiles waste the least), (2) the split-KV workers for the dense
// tail rows, (3) the on-GPU cross-check for the tcgen05 kernel.
//
// Replaces vllm/attention/ascend_trianglemix.py:86-106 (dense matmul / softmax /
// matmul over the full S x S score matrix) of the reference.
#include "tmx_sched.cuh"
#include "tmx_attn.cuh"

namespace tmx {
namespace {

constexpr int kBN = 64;
// head_dim kD is a template parameter of the kernels: 128 (Qwen3) or 64 (the shape of the reference's own
// demo of this op, examples/trianglemix_quickstart.py:163-188; served by this kernel only).

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(sz));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                        uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                          uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0,
                                         uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
// byte offset of 16-byte chunk `chunk` of row `row` in a [rows][kD] bf16 tile, XOR-swizzled
// (kD / 8 = 16 or 8 chunks per row: the XOR with row & 7 stays inside the row either way)
template <int kD>
__device__ __forceinline__ uint32_t swz(int row, int chunk) {
  return (uint32_t)(row * (kD * 2) + ((chunk ^ (row & 7)) << 4));
}

// Load `rows` rows x kD bf16 from global (row stride `stride` elements) into a swizzled tile.
// Rows with index >= valid_rows are zero-filled.
template <int kThreads, int kD>
__device__ __forceinline__ void load_tile(uint32_t sdst, const __nv_bfloat16* g, long long stride,
                                          int rows, int valid_rows, int tid) {
  constexpr int kChunks = kD / 8;
  for (int c = tid; c < rows * kChunks; c += kThreads) {
    int r = c / kChunks, ch = c % kChunks;
    bool ok = r < valid_rows;
    const __nv_bfloat16* src = g + (long long)(ok ? r : 0) * stride + ch * 8;
    cp_async16(sdst + swz<kD>(r, ch), src, ok);
  }
}

template <int BM, int kD>
__global__ void __launch_bounds__(BM * 2, (BM == 64 ? 3 : 1)) attn_mma_kernel(const AttnParams p) {
  constexpr int kThreads = BM * 2;
  constexpr int kTileBytes = kBN * kD * 2;  // 16 KB (kD = 128) / 8 KB (kD = 64)
  constexpr int kChunks = kD / 8;           // 16-byte chunks per row
  constexpr int kKSteps = kD / 16;          // k-steps of the score MMA = 16-column pairs of the output
  extern __shared__ __align__(128) uint8_t smem[];
  // layout: K stage0 | K stage1 | V stage0 | V stage1 ; Q is staged through the V region first
  const uint32_t sK = smem_u32(smem);
  const uint32_t sV = sK + 2 * kTileBytes;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, t4 = lane & 3;

  // ---- which work item? -------------------------------------------------
  const int head = blockIdx.y;
  const int kvh = head / (p.Hq / p.Hkv);
  int seq, n, r0, lo0, hi0, lo1, hi1, part = -1;
  if ((int)blockIdx.x < p.n_slots && p.tail_only) {
    // tail mode: only the query tiles holding dense tail rows (the streaming kernel owns the rest)
    seq = blockIdx.x / p.last_tiles;
    const int ql = blockIdx.x % p.last_tiles;
    n = p.cu[seq + 1] - p.cu[seq];
    const int lrow_ = first_dense_row(n, p.last);
    if (n <= 0 || lrow_ >= n) return;
    r0 = (lrow_ / BM + ql) * BM;
    if (r0 >= n) return;
    const int r1_ = min(n, r0 + BM) - 1;
    if (split_count(r0, r1_, n, kBN, lrow_, p.last, p.mode, BM) > 0) return;   // split workers do it
    const TileRanges t = touched_tiles(r0, r1_, n, kBN, lrow_, p.si, p.sl, p.mode);
    lo0 = t.lo0; hi0 = t.hi0; lo1 = t.lo1; hi1 = t.hi1;
  } else if ((int)blockIdx.x < p.n_slots) {
    const SlotInfo s = p.slots[p.n_slots - 1 - blockIdx.x];   // descending: longest first
    if (s.seq < 0 || s.nsplit > 0) return;
    seq = s.seq; n = s.n; r0 = seq_geom(p, seq).ctx + s.q_tile * BM;   // absolute row (context + chunk)
    lo0 = s.lo0; hi0 = s.hi0; lo1 = s.lo1; hi1 = s.hi1;
  } else {
    // split-KV worker: e -> (sequence, which of the last tiles, chunk)
    int e = blockIdx.x - p.n_slots;
    int chunk = e % p.nsplit_max; e /= p.nsplit_max;
    int ql = e % p.last_tiles; seq = e / p.last_tiles;
    int start = p.cu[seq];
    n = p.cu[seq + 1] - start;
    if (n <= 0) return;
    int lrow_ = first_dense_row(n, p.last);
    int q_tile = lrow_ / BM + ql;
    if (q_tile * BM >= n) return;
    r0 = q_tile * BM;
    int r1_ = min(n, r0 + BM) - 1;
    int ns = split_count(r0, r1_, n, kBN, lrow_, p.last, p.mode, BM);
    if (chunk >= ns) return;
    lo0 = chunk * kSplitTiles; hi0 = min(r1_ / kBN + 1, lo0 + kSplitTiles);
    lo1 = hi1 = 0;
    part = (seq * p.last_tiles + ql) * p.nsplit_max + chunk;
  }
  const SeqGeom gm = seq_geom(p, seq);
  const int seq_start = gm.q_start - gm.ctx;          // query-buffer row of absolute row 0
  const int lrow = first_dense_row(n, p.last);
  int r1 = min(n, r0 + BM) - 1;
  if (p.tail_clip && p.mode == TMX_MODE_TRIANGLE && lrow < n) r1 = min(r1, lrow - 1);
  const int cnt0 = hi0 - lo0, total = cnt0 + (hi1 - lo1);

  const __nv_bfloat16* qg = p.q + (long long)(seq_start + r0) * p.q_stride + head * kD;
  const __nv_bfloat16* kg = p.k + (long long)gm.k_start * p.kv_stride + kvh * kD;
  const __nv_bfloat16* vg = p.v + (long long)gm.k_start * p.kv_stride + kvh * kD;
  const int* pages = p.block_table ? p.block_table + (long long)seq * p.bt_stride : nullptr;

  auto tile_of = [&](int it) { return it < cnt0 ? lo0 + it : lo1 + (it - cnt0); };
  // 64 keys x 128 d of K or V into a swizzled tile; keys >= n are zero-filled.  Paged cache:
  // key j lives at slot pages[j / page_size] * page_size + j % page_size.
  auto load_kv_tile = [&](uint32_t sdst, const __nv_bfloat16* base, int tile) {
    const int valid = min(kBN, n - tile * kBN);
    if (pages == nullptr) {
      load_tile<kThreads, kD>(sdst, base + (long long)tile * kBN * p.kv_stride, p.kv_stride, kBN, valid, tid);
      return;
    }
    for (int c = tid; c < kBN * kChunks; c += kThreads) {
      const int r = c / kChunks, ch = c % kChunks;
      const bool ok = r < valid;
      const int key = tile * kBN + (ok ? r : 0);
      const long long slot = (long long)pages[key / p.page_size] * p.page_size + key % p.page_size;
      cp_async16(sdst + swz<kD>(r, ch), base + slot * p.kv_stride + ch * 8, ok);
    }
  };
  auto issue_kv = [&](int it) {
    const int tile = tile_of(it), st = it & 1;
    load_kv_tile(sK + st * kTileBytes, kg, tile);
    load_kv_tile(sV + st * kTileBytes, vg, tile);
  };

  // ---- prologue: Q through the V region, K tile 0 alongside ------------------
  load_tile<kThreads, kD>(sV, qg, p.q_stride, BM, r1 - r0 + 1, tid);
  load_kv_tile(sK, kg, tile_of(0));
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
  uint32_t qa[kKSteps][4];
  {
    int row = warp * 16 + (lane & 15);
#pragma unroll
    for (int ks = 0; ks < kKSteps; ++ks)
      ldsm_x4(sV + swz<kD>(row, 2 * ks + (lane >> 4)), qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3]);
  }
  __syncthreads();
  load_kv_tile(sV, vg, tile_of(0));
  cp_async_commit();

  float o[kChunks][4];
#pragma unroll
  for (int i = 0; i < kChunks; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  const float sc = p.scale_log2;
  const int row_a = r0 + warp * 16 + g, row_b = row_a + 8;   // rows inside the sequence

  for (int it = 0; it < total; ++it) {
    const int st = it & 1;
    if (it + 1 < total) issue_kv(it + 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();

    const int tile = tile_of(it);
    const int c0 = tile * kBN;
    const uint32_t kbase = sK + st * kTileBytes, vbase = sV + st * kTileBytes;

    // ---- S = Q K^T ---------------------------------------------------------
    float s[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < kKSteps; ++ks) {
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        int key = np * 16 + ((lane >> 4) << 3) + (lane & 7);
        int chunk = 2 * ks + ((lane >> 3) & 1);
        uint32_t b0, b1, b2, b3;
        ldsm_x4(kbase + swz<kD>(key, chunk), b0, b1, b2, b3);
        mma_bf16(s[2 * np], qa[ks], b0, b1);
        mma_bf16(s[2 * np + 1], qa[ks], b2, b3);
      }
    }

    // ---- mask (PARTIAL tiles only) + scale -------------------------------------
    const bool full = p.bitmask == nullptr &&
                      tile_is_full(r0, r1, c0, c0 + kBN - 1, n, lrow, p.si, p.sl, p.mode);
    if (full) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s[i][0] *= sc; s[i][1] *= sc; s[i][2] *= sc; s[i][3] *= sc;
      }
    } else if (p.bitmask == nullptr) {  // analytic predicate
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        int j = c0 + i * 8 + t4 * 2;
        s[i][0] = attends(row_a, j, n, lrow, p.si, p.sl, p.mode) ? s[i][0] * sc : -INFINITY;
        s[i][1] = attends(row_a, j + 1, n, lrow, p.si, p.sl, p.mode) ? s[i][1] * sc : -INFINITY;
        s[i][2] = attends(row_b, j, n, lrow, p.si, p.sl, p.mode) ? s[i][2] * sc : -INFINITY;
        s[i][3] = attends(row_b, j + 1, n, lrow, p.si, p.sl, p.mode) ? s[i][3] * sc : -INFINITY;
      }
    } else {
      // bitmask source: 64 columns = 2 words per row
      const uint32_t* wa = p.bitmask + (long long)min(row_a, n - 1) * p.bitmask_stride + (c0 >> 5);
      const uint32_t* wb = p.bitmask + (long long)min(row_b, n - 1) * p.bitmask_stride + (c0 >> 5);
      uint32_t a0 = wa[0], a1 = wa[1], b0 = wb[0], b1 = wb[1];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        int jj = (i & 3) * 8 + t4 * 2;
        uint32_t wa_ = i < 4 ? a0 : a1, wb_ = i < 4 ? b0 : b1;
        s[i][0] = ((wa_ >> jj) & 1u) ? s[i][0] * sc : -INFINITY;
        s[i][1] = ((wa_ >> (jj + 1)) & 1u) ? s[i][1] * sc : -INFINITY;
        s[i][2] = ((wb_ >> jj) & 1u) ? s[i][2] * sc : -INFINITY;
        s[i][3] = ((wb_ >> (jj + 1)) & 1u) ? s[i][3] * sc : -INFINITY;
      }
    }

    // ---- online softmax ------------------------------------------------------------
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      mx[0] = fmaxf(mx[0], fmaxf(s[i][0], s[i][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[i][2], s[i][3]));
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
    }
    float alpha[2], muse[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      float m_new = fmaxf(m_run[r], mx[r]);
      muse[r] = (m_new == -INFINITY) ? 0.f : m_new;       // row fully masked so far
      alpha[r] = fast_exp2(m_run[r] - muse[r]);               // m_run == -inf -> 0
      m_run[r] = m_new;
    }
    float rs[2] = {0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      s[i][0] = fast_exp2(s[i][0] - muse[0]); s[i][1] = fast_exp2(s[i][1] - muse[0]);
      s[i][2] = fast_exp2(s[i][2] - muse[1]); s[i][3] = fast_exp2(s[i][3] - muse[1]);
      rs[0] += s[i][0] + s[i][1];
      rs[1] += s[i][2] + s[i][3];
    }
    l_run[0] = l_run[0] * alpha[0] + rs[0];
    l_run[1] = l_run[1] * alpha[1] + rs[1];
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      o[i][0] *= alpha[0]; o[i][1] *= alpha[0]; o[i][2] *= alpha[1]; o[i][3] *= alpha[1];
    }

    // ---- O += P V ------------------------------------------------------------------
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_bf16(s[2 * kk][0], s[2 * kk][1]);
      pa[1] = pack_bf16(s[2 * kk][2], s[2 * kk][3]);
      pa[2] = pack_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[3] = pack_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
      for (int dp = 0; dp < kKSteps; ++dp) {
        int key = kk * 16 + (((lane >> 3) & 1) << 3) + (lane & 7);
        int chunk = 2 * dp + (lane >> 4);
        uint32_t b0, b1, b2, b3;
        ldsm_x4_t(vbase + swz<kD>(key, chunk), b0, b1, b2, b3);
        mma_bf16(o[2 * dp], pa, b0, b1);
        mma_bf16(o[2 * dp + 1], pa, b2, b3);
      }
    }
    __syncthreads();   // stage `st` may be overwritten by the load issued next iteration
  }
  cp_async_wait<0>();

  // ---- epilogue ------------------------------------------------------------------------
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  if (part >= 0) {
    // fp32 partial: O (unnormalised), m (log2 domain), l
    float* po = p.part_o + ((long long)part * p.Hq + head) * BM * kD;
    float2* pml = p.part_ml + ((long long)part * p.Hq + head) * BM;
    int ra = warp * 16 + g, rb = ra + 8;
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      *reinterpret_cast<float2*>(po + ra * kD + i * 8 + t4 * 2) = make_float2(o[i][0], o[i][1]);
      *reinterpret_cast<float2*>(po + rb * kD + i * 8 + t4 * 2) = make_float2(o[i][2], o[i][3]);
    }
    if (t4 == 0) {
      pml[ra] = make_float2(m_run[0], l_run[0]);
      pml[rb] = make_float2(m_run[1], l_run[1]);
    }
    return;
  }
  float inv[2] = {l_run[0] > 0.f ? fast_rcp(l_run[0]) : 0.f, l_run[1] > 0.f ? fast_rcp(l_run[1]) : 0.f};
  // stage the bf16 tile through shared memory (K region is free now) for 16-byte row stores
  {
    int ra = warp * 16 + g, rb = ra + 8;
#pragma unroll
    for (int i = 0; i < kChunks; ++i) {
      uint32_t va = pack_bf16(o[i][0] * inv[0], o[i][1] * inv[0]);
      uint32_t vb = pack_bf16(o[i][2] * inv[1], o[i][3] * inv[1]);
      asm volatile("st.shared.b32 [%0], %1;\n" ::"r"(sK + swz<kD>(ra, i) + t4 * 4), "r"(va));
      asm volatile("st.shared.b32 [%0], %1;\n" ::"r"(sK + swz<kD>(rb, i) + t4 * 4), "r"(vb));
    }
  }
  __syncthreads();
  __nv_bfloat16* og = p.out + (long long)(seq_start + r0) * p.o_stride + head * kD;
  const int valid = r1 - r0 + 1;
  for (int c = tid; c < BM * kChunks; c += kThreads) {
    int r = c / kChunks, ch = c % kChunks;
    if (r < valid) {
      uint4 v;
      asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];\n"
                   : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(sK + swz<kD>(r, ch)));
      *reinterpret_cast<uint4*>(og + (long long)r * p.o_stride + ch * 8) = v;
    }
  }
  if (p.lse != nullptr && t4 == 0) {
    if (row_a <= r1) p.lse[(long long)head * p.total_T + seq_start + row_a] =
        m_run[0] * 0.6931471805599453f + logf(l_run[0]);
    if (row_b <= r1) p.lse[(long long)head * p.total_T + seq_start + row_b] =
        m_run[1] * 0.6931471805599453f + logf(l_run[1]);
  }
}

// Merge the split-KV partials of one (sequence, last tile, head): a CTA takes 8 rows, one thread
// per (row, 4 columns); grid.x = B * last_tiles * BM / 8 so that the merge is a bandwidth-bound
// pass over the partials instead of 64 serial CTAs.
template <int BM, int kD>
__global__ void __launch_bounds__(256) combine_kernel(const AttnParams p) {
  constexpr int kRowsPerCta = 8;
  int e = blockIdx.x / (BM / kRowsPerCta);
  const int rblk = blockIdx.x % (BM / kRowsPerCta);
  int ql = e % p.last_tiles, seq = e / p.last_tiles;
  int head = blockIdx.y;
  int start = p.cu[seq], n = p.cu[seq + 1] - start;
  if (n <= 0) return;
  int lrow = first_dense_row(n, p.last);
  int q_tile = lrow / BM + ql;
  int r0 = q_tile * BM;
  if (r0 >= n) return;
  int r1 = min(n, r0 + BM) - 1;
  int ns = split_count(r0, r1, n, kBN, lrow, p.last, p.mode, BM);
  if (ns == 0) return;
  long long base = (long long)(seq * p.last_tiles + ql) * p.nsplit_max;
  const int r = rblk * kRowsPerCta + (threadIdx.x >> 5), c4 = (threadIdx.x & 31) * 4;
  if (r0 + r > r1 || c4 >= kD) return;     // kD = 64: the upper half-warp has no columns
  float mmax = -INFINITY;
  for (int s = 0; s < ns; ++s)
    mmax = fmaxf(mmax, p.part_ml[((base + s) * p.Hq + head) * BM + r].x);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  float l = 0.f;
#pragma unroll 4
  for (int s = 0; s < ns; ++s) {
    float2 ml = p.part_ml[((base + s) * p.Hq + head) * BM + r];
    float w = (ml.x == -INFINITY) ? 0.f : fast_exp2(ml.x - mmax);
    l += ml.y * w;
    float4 v = *reinterpret_cast<const float4*>(p.part_o + (((base + s) * p.Hq + head) * BM + r) * kD + c4);
    acc.x += v.x * w; acc.y += v.y * w; acc.z += v.z * w; acc.w += v.w * w;
  }
  float inv = l > 0.f ? 1.f / l : 0.f;
  __nv_bfloat16* og = p.out + (long long)(start + r0 + r) * p.o_stride + head * kD + c4;
  uint2 packed = make_uint2(pack_bf16(acc.x * inv, acc.y * inv), pack_bf16(acc.z * inv, acc.w * inv));
  *reinterpret_cast<uint2*>(og) = packed;
  if (p.lse != nullptr && c4 == 0)
    p.lse[(long long)head * p.total_T + start + r0 + r] = mmax * 0.6931471805599453f + logf(l);
}

}  // namespace

template <int BM, int kD>
int launch_attn_mma_t(const AttnParams& p, cudaStream_t s) {
  constexpr int smem = 4 * kBN * kD * 2;        // K and V, two stages each: 64 KB (kD = 128) / 32 KB (kD = 64)
  if constexpr (smem > 48 * 1024)               // (kD = 64 stays under the 48 KB every kernel may use)
    TMX_SET_SMEM_ATTR((attn_mma_kernel<BM, kD>), BM == 64 ? kAttrMma64 : kAttrMma128, smem);
  int extra = (p.last_tiles > 0 && p.nsplit_max > 0) ? p.B * p.last_tiles * p.nsplit_max : 0;
  dim3 grid((unsigned)(p.n_slots + extra), (unsigned)p.Hq);
  attn_mma_kernel<BM, kD><<<grid, BM * 2, smem, s>>>(p);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  if (extra > 0) {
    dim3 cg((unsigned)(p.B * p.last_tiles * (BM / 8)), (unsigned)p.Hq);
    combine_kernel<BM, kD><<<cg, 256, 0, s>>>(p);
    count_launch();
    TMX_CUDA_TRY(cudaGetLastError());
  }
  return TMX_OK;
}

int launch_attn_mma(const AttnParams& p, int bm, cudaStream_t s) {
  if (p.D == 64) return bm == 64 ? launch_attn_mma_t<64, 64>(p, s) : launch_attn_mma_t<128, 64>(p, s);
  return bm == 64 ? launch_attn_mma_t<64, 128>(p, s) : launch_attn_mma_t<128, 128>(p, s);
}

}  // namespace tmx

// Parameters shared by the attention kernels and their launchers.
#pragma once
#include "tmx_sched.cuh"

namespace tmx {

struct AttnParams {
  const __nv_bfloat16* q;
  const __nv_bfloat16* k;     // contiguous K/V [sum kv_len, Hkv, D], or the paged cache base
  const __nv_bfloat16* v;
  __nv_bfloat16* out;
  float* lse;                 // [Hq, total_T] or null
  const int* cu;              // [B + 1] query offsets
  // ---- context (chunked prefill / decode) --------------------------
  // kv_lens[b] = total keys of sequence b (context + this call's query rows); null: == query length.
  // Query row r of the call is row (kv_len - q_len + r) of the mask.
  const int* kv_lens;
  const int* cu_k;            // [B + 1] key offsets of a contiguous K/V buffer; null: == cu
  const int* block_table;     // [B, bt_stride] page ids of a paged cache; null: contiguous K/V
  long long bt_stride;
  int page_size;              // tokens per page (paged cache)
  const SlotInfo* slots;      // [n_slots]
  const uint32_t* bitmask;    // optional packed mask (B == 1)
  long long bitmask_stride;   // words per row
  float* part_o;              // split-KV partials
  float2* part_ml;
  long long q_stride, kv_stride, o_stride;   // elements per token (paged: per cache slot)
  int B, total_T, max_seqlen, Hq, Hkv;
  int D;                      // head_dim: 128 (every kernel) or 64 (mma.sync kernel only)
  long long total_kv;         // rows of the K/V buffer (contiguous: tokens; paged: num_blocks * page_size)
  int n_slots, last_tiles, nsplit_max;
  int si, sl, last, mode;
  int tail_only;              // mma kernel: blockIdx.x < n_slots enumerates (sequence, tail tile) only
  int tail_bm;                // stream kernel: query-tile size of the dense-tail path
  int tail_clip;              // block-index kernels: rows >= the first dense row belong to the tail path
  float scale_log2;           // scale * log2(e)
  long long* trace;           // developer build only (tmx_dev_set_trace): clock64 timeline buffer, else null
};

// Per-sequence geometry, identical in every kernel.
struct SeqGeom {
  int q_start, nq, n, ctx, k_start;
};
__device__ __forceinline__ SeqGeom seq_geom(const AttnParams& p, int seq) {
  SeqGeom g;
  g.q_start = p.cu[seq];
  g.nq = p.cu[seq + 1] - g.q_start;
  g.n = p.kv_lens ? p.kv_lens[seq] : g.nq;
  g.ctx = g.n - g.nq;
  g.k_start = p.block_table ? 0 : (p.cu_k ? p.cu_k[seq] : g.q_start);
  return g;
}

int build_slots(SlotInfo* slots, int n_slots, const AttnParams& p, int bm, int bn, cudaStream_t s);
int launch_attn_mma(const AttnParams& p, int bm, cudaStream_t s);
// streaming part of triangle layers (tmx_attn_stream3.cu)
bool stream_supported(int Hq, int Hkv, int si, int sl);
int stream_block_m(int Hq, int Hkv);
int launch_attn_stream3(const AttnParams& p, cudaStream_t s);   // persistent, pipelined
int launch_attn_sm100_v3(const AttnParams& p, int variant, cudaStream_t s);   // 128-key tiles
// CTA-pair kernel (cta_group::2): GQA ratios that are multiples of 4 (tmx_attn_sm100_pair.cu)
bool pair_supported(const AttnParams& p);
int launch_attn_sm100_pair(const AttnParams& p, cudaStream_t s);
// dense tail rows of triangle layers on the tensor cores (split-KV + merge, tmx_attn_sm100_v3.cu)
bool tail_tc_supported(int Hq, int Hkv, float scale_log2);
int tail_tc_block_m(int Hq, int Hkv);
int launch_attn_tail_tc(const AttnParams& p, cudaStream_t s);
// Developer switch (TMX_SM100_VARIANT): read ONLY by a -DTMX_DEV build; the product library ignores it.
int sm100_variant();

}  // namespace tmx

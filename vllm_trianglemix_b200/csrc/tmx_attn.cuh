// Parameters shared by the attention kernels and their launchers.
#pragma once
#include "tmx_sched.cuh"

namespace tmx {

struct AttnParams {
  const __nv_bfloat16* q;
  const __nv_bfloat16* k;
  const __nv_bfloat16* v;
  __nv_bfloat16* out;
  float* lse;                 // [Hq, total_T] or null
  const int* cu;              // [B + 1]
  const SlotInfo* slots;      // [n_slots]
  const uint32_t* bitmask;    // optional packed mask (B == 1)
  long long bitmask_stride;   // words per row
  float* part_o;              // split-KV partials
  float2* part_ml;
  long long q_stride, kv_stride, o_stride;   // elements per token
  int B, total_T, max_seqlen, Hq, Hkv;
  int n_slots, last_tiles, nsplit_max;
  int si, sl, last, mode;
  int tail_only;              // mma kernel: blockIdx.x < n_slots enumerates (sequence, tail tile) only
  int tail_bm;                // stream kernel: query-tile size of the dense-tail path
  float scale_log2;           // scale * log2(e)
};

int build_slots(SlotInfo* slots, int n_slots, const int* cu, int B, int max_seqlen, int bm, int bn,
                int si, int sl, int last, int mode, cudaStream_t s);
int launch_attn_mma(const AttnParams& p, int bm, cudaStream_t s);
// streaming part of triangle layers (tmx_attn_stream.cu)
bool stream_supported(int Hq, int Hkv, int si, int sl);
int stream_block_m(int Hq, int Hkv);
int launch_attn_stream(const AttnParams& p, cudaStream_t s);    // one CTA per tile (v1)
int launch_attn_stream2(const AttnParams& p, cudaStream_t s);   // persistent, pipelined (default)
int launch_attn_sm100(const AttnParams& p, int variant, cudaStream_t s);      // 64-key tiles (v2)
int launch_attn_sm100_v3(const AttnParams& p, int variant, cudaStream_t s);   // 128-key tiles (v3, default)
// dense tail rows of triangle layers on the tensor cores (split-KV + merge, tmx_attn_sm100_v3.cu)
bool tail_tc_supported(int Hq, int Hkv, float scale_log2);
int tail_tc_block_m(int Hq, int Hkv);
int launch_attn_tail_tc(const AttnParams& p, cudaStream_t s);
// TMX_SM100_VARIANT env (developer switch), default 0: bit 7 selects the 64-key kernel, the low
// bits are that kernel's own timing/emulation switches.
int sm100_variant();
inline bool sm100_use_v2(int variant) { return (variant & 128) != 0; }
inline int sm100_block_n(int variant) { return sm100_use_v2(variant) ? 64 : 128; }

}  // namespace tmx

// Schedule shared by the block-index kernel and the attention kernels.
//
// A "slot" is one query tile of block_m rows of one sequence.  Slot ids are
// assigned without a prefix scan:  slot(b, q) = cu_seqlens[b] / block_m + b + q,
// which is strictly increasing in (b, q) and bounded by total_T / block_m + B.
// The block-index kernel writes one SlotInfo per slot; the attention kernels read
// it back (slot -> sequence, tile, touched KV tile ranges).
//
// Triangle layers: the `last - 1` dense rows at the end of every sequence attend
// all earlier keys -- 2/3 of a 128K triangle layer's work sits in ONE query tile.
// Those tiles are split along KV into chunks of kSplitTiles tiles, each chunk a
// CTA of its own writing an fp32 partial (O, m, l), merged by a combine kernel.
#pragma once
#include "tmx_common.cuh"

namespace tmx {

struct __align__(16) SlotInfo {
  int seq;      // -1: unused slot
  int q_tile;   // tile index inside the sequence
  int lo0, hi0; // first touched KV tile range [lo0, hi0)
  int lo1, hi1; // second range (empty when lo1 == hi1)
  int nsplit;   // 0: handled by the slot's own CTA; >0: split into nsplit chunk CTAs
  int n;        // sequence length
};

constexpr int kSplitTiles = 32;       // KV tiles per split chunk
constexpr int kSplitMinTiles = 64;    // split only tiles touching at least this many KV tiles
constexpr int kSplitMaxLastRows = 512;    // mma.sync split path (fp32 partials grow with last x tiles)
constexpr int kTailTcMaxRows = 8192;      // tcgen05 tail path: `last - 1` dense rows it will split

__host__ __device__ __forceinline__ int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Number of query tiles per sequence that can hold "last" rows (upper bound, host-known).
__host__ __device__ __forceinline__ int last_tiles_max(int last, int bm) {
  int rows = last - 1;
  if (rows <= 0 || rows > kSplitMaxLastRows) return 0;
  return ceil_div(rows, bm) + 1;
}
__host__ __device__ __forceinline__ int nsplit_max(int max_seqlen, int bn) {
  return ceil_div(ceil_div(max_seqlen, bn), kSplitTiles);
}

// tcgen05 tail path: a tail tile touching `tiles` 128-key KV tiles is cut into this many
// contiguous chunks (at least kTailMinTiles tiles each, at most the plan's nsplit_max).
constexpr int kTailMinTiles = 4;
__host__ __device__ __forceinline__ int tail_nsplit(int tiles, int nsplit_max) {
  int ns = tiles / kTailMinTiles;
  ns = ns < 1 ? 1 : ns;
  return ns < nsplit_max ? ns : nsplit_max;
}

// Does query tile q (rows [r0, r1]) of a length-n sequence get split?  Returns nsplit or 0.
__host__ __device__ __forceinline__ int split_count(int r0, int r1, int n, int bn, int lrow,
                                                    int last, int mode, int bm) {
  if (mode != TMX_MODE_TRIANGLE || last_tiles_max(last, bm) == 0) return 0;
  if (r1 < lrow) return 0;
  int tiles = r1 / bn + 1;
  if (tiles < kSplitMinTiles) return 0;
  return ceil_div(tiles, kSplitTiles);
}

}  // namespace tmx

// Block-index kernel: per (sequence, query tile) the touched KV tile ranges, in
// closed form -- O(N / block_m) work, nothing N x N is ever materialised.
// The idea of a per-row-block list of KV blocks is the one the reference tree's
// csrc/attention/vertical_slash_index.cu uses for MInference; the TriangleMix
// pattern (vertical sink + slash window + dense tail) needs at most two ranges.
#include "tmx_attn.cuh"

namespace tmx {

__global__ void build_slots_kernel(SlotInfo* __restrict__ slots, const AttnParams p, int bm, int bn) {
  const int b = blockIdx.y;
  const SeqGeom g = seq_geom(p, b);
  if (g.nq <= 0 || g.ctx < 0) return;
  const int ntiles = ceil_div(g.nq, bm);
  const int lrow = first_dense_row(g.n, p.last);
  const int base = g.q_start / bm + b;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < ntiles; q += gridDim.x * blockDim.x) {
    // absolute rows of the tile inside the (context + chunk) sequence
    const int r0 = g.ctx + q * bm;
    int r1 = g.ctx + min(g.nq, (q + 1) * bm) - 1;
    SlotInfo s;
    s.seq = b; s.q_tile = q; s.n = g.n;
    if (p.tail_clip && p.mode == TMX_MODE_TRIANGLE && lrow < g.n) {
      // the dense tail rows [lrow, n) are served by the split-KV tail path
      if (r0 >= lrow) { s.seq = -1; r1 = r0; }
      else r1 = min(r1, lrow - 1);
    }
    const TileRanges t = touched_tiles(r0, r1, g.n, bn, lrow, p.si, p.sl, p.mode);
    s.lo0 = t.lo0; s.hi0 = t.hi0; s.lo1 = t.lo1; s.hi1 = t.hi1;
    s.nsplit = (g.ctx == 0 && !p.tail_clip) ? split_count(r0, r1, g.n, bn, lrow, p.last, p.mode, bm) : 0;
    slots[base + q] = s;
  }
}

int build_slots(SlotInfo* slots, int n_slots, const AttnParams& p, int bm, int bn, cudaStream_t s) {
  TMX_CUDA_TRY(cudaMemsetAsync(slots, 0xff, sizeof(SlotInfo) * (size_t)n_slots, s));
  int tiles = ceil_div(p.max_seqlen > 0 ? p.max_seqlen : 1, bm);
  dim3 grid((unsigned)min(ceil_div(tiles, 128), 1024), (unsigned)p.B);
  build_slots_kernel<<<grid, 128, 0, s>>>(slots, p, bm, bn);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

namespace {
__global__ void export_index_kernel(int* __restrict__ ranges, int* __restrict__ counts,
                                    int* __restrict__ full, const int* __restrict__ cu, int n_slots,
                                    int bm, int bn, int si, int sl, int last, int mode) {
  // pass 1: clear (grid-stride over slots), pass 2 is fused: each (b, q) thread overwrites its slot
  int b = blockIdx.y;
  int start = cu[b], n = cu[b + 1] - start;
  if (n <= 0) return;
  int ntiles = ceil_div(n, bm);
  int lrow = first_dense_row(n, last);
  int base = start / bm + b;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < ntiles; q += gridDim.x * blockDim.x) {
    int r0 = q * bm, r1 = min(n, r0 + bm) - 1;
    TileRanges t = touched_tiles(r0, r1, n, bn, lrow, si, sl, mode);
    int slot = base + q;
    ranges[slot * 4 + 0] = t.lo0; ranges[slot * 4 + 1] = t.hi0;
    ranges[slot * 4 + 2] = t.lo1; ranges[slot * 4 + 3] = t.hi1;
    counts[slot] = (t.hi0 - t.lo0) + (t.hi1 - t.lo1);
    int f0 = 0, f1 = 0;
    for (int j = t.lo0; j < t.hi0; ++j)
      f0 += tile_is_full(r0, r1, j * bn, j * bn + bn - 1, n, lrow, si, sl, mode);
    for (int j = t.lo1; j < t.hi1; ++j)
      f1 += tile_is_full(r0, r1, j * bn, j * bn + bn - 1, n, lrow, si, sl, mode);
    full[slot * 2 + 0] = f0; full[slot * 2 + 1] = f1;
  }
}
}  // namespace
}  // namespace tmx

using namespace tmx;

extern "C" int tmx_block_index(int32_t* ranges, int32_t* counts, int32_t* full,
                               const int32_t* cu_seqlens, int B, int total_T, int block_m,
                               int block_n, int si, int sl, int last, int mode,
                               tmx_stream_t stream) {
  if (B < 0 || total_T < 0 || block_m <= 0 || block_n <= 0) return TMX_ERR_BAD_ARG;
  if (mode != TMX_MODE_DENSE_CAUSAL && mode != TMX_MODE_TRIANGLE) return TMX_ERR_BAD_ARG;
  if (B == 0) return TMX_OK;
  if (!ranges || !counts || !full || !cu_seqlens) return TMX_ERR_BAD_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  int n_slots = total_T / block_m + B;
  TMX_CUDA_TRY(cudaMemsetAsync(ranges, 0, sizeof(int) * 4 * (size_t)n_slots, s));
  TMX_CUDA_TRY(cudaMemsetAsync(counts, 0, sizeof(int) * (size_t)n_slots, s));
  TMX_CUDA_TRY(cudaMemsetAsync(full, 0, sizeof(int) * 2 * (size_t)n_slots, s));
  int tiles = ceil_div(total_T > 0 ? total_T : 1, block_m);
  dim3 grid((unsigned)min(ceil_div(tiles, 128), 1024), (unsigned)B);
  export_index_kernel<<<grid, 128, 0, s>>>(ranges, counts, full, cu_seqlens, n_slots, block_m,
                                           block_n, si, sl, last, mode);
  count_launch();
  TMX_CUDA_TRY(cudaGetLastError());
  return TMX_OK;
}

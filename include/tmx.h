/*
 * tmx.h -- C ABI of the B200-native TriangleMix prefill path (libtmx.so).
 *
 * The reference (ackerm0n/vllm-trianglemix) has no native ABI for this path: it is
 * 1.1k lines of eager PyTorch.  Each entry point below therefore cites the Python
 * function whose work it replaces (paths relative to the reference root).  The
 * Python host (vllm_trianglemix_b200/) keeps the reference's class and method
 * signatures and calls these through ctypes with raw device pointers.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - the caller owns all buffers (outputs and workspace); the library allocates
 *     nothing persistent except cached TMA descriptors (per thread) and per-DEVICE
 *     kernel attributes / SM count, so one process may drive several GPUs;
 *   - the library reads no environment variable (developer timing variants exist
 *     only in a `make DEV=1` build);
 *   - every call only ENQUEUES work on `stream` (a cudaStream_t) -- no host sync;
 *   - return value 0 = TMX_OK, otherwise a tmx_status (see tmx_strerror);
 *   - mask parameters: si = num_sink_tokens, sl = sliding_window_size,
 *     last = num_last_tokens  (vllm/attention/trianglemix_attention.py:14-36);
 *   - element predicate for query row i, key column j of a length-N sequence
 *       triangle: j <= i && (j <= si || i - j <= sl || N - i < last)
 *       dense   : j <= i
 *     which is bit-exact with the reference's loops (:82-89, :126-129, :192-196).
 */
#ifndef TMX_H_
#define TMX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* tmx_stream_t; /* == cudaStream_t */

typedef enum {
  TMX_OK = 0,
  TMX_ERR_BAD_ARG = 1,      /* null pointer, negative size, unknown enum        */
  TMX_ERR_BAD_SHAPE = 2,    /* head_dim not 128 / 64, Hq % Hkv != 0, ...        */
  TMX_ERR_ALIGNMENT = 3,    /* pointer / stride not 16-byte aligned             */
  TMX_ERR_WORKSPACE = 4,    /* workspace null or smaller than tmx_attn_workspace_bytes */
  TMX_ERR_ARCH = 5,         /* device is not compute capability 10.x            */
  TMX_ERR_CUDA = 6,         /* a CUDA runtime/driver call failed (see tmx_last_cuda_error) */
  TMX_ERR_UNSUPPORTED = 7,  /* valid request this build cannot serve            */
  TMX_ERR_STATIC_SCRATCH = 8 /* tmx_prefill_attn_static: the request needs more scratch than
                                TMX_STATIC_SCRATCH_BYTES -- use tmx_prefill_attn with a caller workspace */
} tmx_status;

/* mask_mode of a layer: the output of should_use_triangle()
 * (vllm/attention/trianglemix_attention.py:263-278). */
enum { TMX_MODE_DENSE_CAUSAL = 0, TMX_MODE_TRIANGLE = 1,
       /* no mask at all (bidirectional): what the reference op computes when neither the
        * helper nor the caller supplies a mask (vllm/attention/ascend_trianglemix.py:96-97,
        * e.g. every seq_len <= 2048).  Attention only; not a mask-generator mode. */
       TMX_MODE_NONE = 2 };

/* the five 0/1 pieces the generator exposes (:53, :93, :133, :173, :200) */
enum { TMX_PIECE_STREAMING = 0, TMX_PIECE_LAST = 1, TMX_PIECE_MIDDLE = 2,
       TMX_PIECE_TRIANGLE = 3, TMX_PIECE_DENSE = 4 };

/* element types of the compat (N x N) mask outputs */
enum { TMX_F32 = 0, TMX_BF16 = 1, TMX_F16 = 2 };

/* which attention kernel serves tmx_prefill_attn */
enum { TMX_KERNEL_AUTO = 0,   /* tcgen05 kernel where it applies, else mma.sync */
       TMX_KERNEL_MMA = 1,    /* warp-level mma.sync kernel (also the on-GPU cross-check) */
       TMX_KERNEL_TCGEN05 = 2, /* tcgen05/TMEM/TMA kernel, one CTA per two query heads */
       TMX_KERNEL_TCGEN05_PAIR = 3 /* tcgen05 on CTA pairs (cta_group::2), four query heads of a KV head
                                      per cluster; needs Hq/Hkv % 4 == 0, else TMX_ERR_UNSUPPORTED */ };

const char* tmx_strerror(int status);
const char* tmx_last_cuda_error(void);
int tmx_version(void);

/* ---- masks -------------------------------------------------------------- */

/* Packed bitmask of one piece.  Row i occupies row_stride_words uint32 words;
 * bit b (LSB first) of word w = piece[i, 32*w + b]; bits at columns >= N and
 * padding words are 0.  row_stride_words must be a multiple of 4 (16-byte rows)
 * and >= ceil(N/32).  Replaces the O(N^2) Python loops
 * _create_{streaming,last,middle,triangle,dense}_mask
 * (vllm/attention/trianglemix_attention.py:53-219) for native consumers. */
int tmx_mask_bits(uint32_t* out, int64_t row_stride_words, int N, int si, int sl,
                  int last, int piece, tmx_stream_t stream);

/* The same pieces as a dense N x N tensor of 0/1 in `dtype` with leading
 * dimension ld (elements) -- the exact return value of the reference's
 * _create_*_mask methods (:53-219). */
int tmx_mask_dense01(void* out, int dtype, int64_t ld, int N, int si, int sl,
                     int last, int piece, tmx_stream_t stream);

/* Additive form {0, -inf} of the triangle / dense-causal mask in `dtype`:
 * get_attention_mask's return value (:221-261, the torch.where at :254-258). */
int tmx_mask_additive(void* out, int dtype, int64_t ld, int N, int si, int sl,
                      int last, int mode, tmx_stream_t stream);

/* `-inf -> fill` rewrite of an additive mask, in place or out of place:
 * AscendNPUTriangleMixOptimizer.optimize_mask_for_npu
 * (vllm/attention/ascend_trianglemix.py:114-141). numel elements of `dtype`. */
int tmx_mask_rewrite_neg_inf(void* out, const void* in, int dtype, int64_t numel,
                             float fill, tmx_stream_t stream);

/* Arbitrary additive N x N mask (values 0 / -inf, as a caller may pass in
 * `attn_mask`, vllm/attention/ascend_trianglemix.py:57,84,97) -> packed bitmask
 * (warp ballot).  0 = attend; -inf or any value <= -1e4 = masked (covers the -1e6
 * that optimize_mask_for_npu writes, :135-139: its softmax weight is exactly 0 in
 * fp32 next to any attended key).  *bad_count (device int32, may be null) receives
 * the number of elements that are neither. */
int tmx_mask_bits_from_additive(uint32_t* out, int64_t row_stride_words,
                                const void* in, int dtype, int64_t ld, int N,
                                int32_t* bad_count, tmx_stream_t stream);

/* ---- block index ---------------------------------------------------------
 * Per (sequence, query tile of block_m rows) the KV tiles (block_n columns) the
 * mask touches, as at most two half-open tile ranges, plus the tile classes.
 * ranges[t*4 + {0,1,2,3}] = lo0, hi0, lo1, hi1 (hi exclusive; empty range lo==hi);
 * counts[t] = number of touched KV tiles; full[t*2+{0,1}] = number of leading
 * tiles of range0 / range1 that are FULL (no per-element predicate needed) --
 * remaining touched tiles are PARTIAL, everything else is SKIPPED.
 * Query tile slot of (sequence b, tile q) = cu_seqlens[b] / block_m + b + q;
 * n_slots = total_T / block_m + B slots must be provided; unused slots get
 * empty ranges.  cu_seqlens: int32 [B+1] on the device. */
int tmx_block_index(int32_t* ranges, int32_t* counts, int32_t* full,
                    const int32_t* cu_seqlens, int B, int total_T, int block_m,
                    int block_n, int si, int sl, int last, int mode,
                    tmx_stream_t stream);

/* ---- masked GQA prefill attention ------------------------------------------
 * out[t, h, :] = softmax_j( q[t,h,:] . k[j, h / (Hq/Hkv), :] * scale + mask[t,j] ) v[j, ...]
 * over the keys j of t's own sequence, i.e. AscendTriangleMixAttention.forward
 * (vllm/attention/ascend_trianglemix.py:86-106) with the additive mask of
 * get_attention_mask folded into tile skipping + an in-register predicate, in the
 * varlen layout of vLLM's AttentionImpl.forward
 * (vllm/attention/backends/abstract.py:158-171; q [T,Hq,D], k/v [T,Hkv,D]).
 * bf16 in, bf16 out, fp32 accumulation.  D must be 128 (Qwen3; every kernel) or 64 (the shape of the
 * reference's own demo of this op, examples/trianglemix_quickstart.py:163-188; served by the mma.sync kernel:
 * TMX_KERNEL_AUTO picks it, an explicit tcgen05 request returns TMX_ERR_UNSUPPORTED).
 * strides are in elements between consecutive tokens; heads are contiguous
 * (stride D).  cu_seqlens: int32 [B+1] device.  lse (optional): float [Hq, total_T],
 * natural-log-sum-exp of the scaled masked scores.
 * bitmask (optional): packed [max_seqlen, stride] mask in tmx_mask_bits layout,
 * indexed by the row/column INSIDE each sequence (every sequence shares it, as the
 * reference broadcasts one (S,S) mask over the batch).  When given, every touched
 * tile reads mask bits instead of evaluating the predicate; `mode` then only
 * selects which tiles are touched (TMX_MODE_NONE = all).  Used for caller-supplied
 * masks and as a cross-check of the analytic path. */
size_t tmx_attn_workspace_bytes(int B, int total_T, int max_seqlen, int Hq, int Hkv,
                                int D, int si, int sl, int last, int mode, int kernel);

int tmx_prefill_attn(void* out, const void* q, const void* k, const void* v,
                     const int32_t* cu_seqlens, int B, int total_T, int max_seqlen,
                     int Hq, int Hkv, int D, int64_t q_stride_t, int64_t kv_stride_t,
                     int64_t o_stride_t, float scale, int si, int sl, int last,
                     int mode, const uint32_t* bitmask, int64_t bitmask_stride_words,
                     float* lse, void* workspace, size_t workspace_bytes, int kernel,
                     tmx_stream_t stream);

/* ---- allocation-free variant of tmx_prefill_attn ---------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, checks, kernels and results as tmx_prefill_attn, minus (workspace, workspace_bytes): the scratch
 * (block index + split-KV partials) is a file-scope `static __device__` array of TMX_STATIC_SCRATCH_BYTES inside the
 * library -- every GPU that loads libtmx.so holds those bytes until the process exits.  The call allocates nothing:
 * no cudaMalloc / cudaFree / allocating library routine; its one-time set-up per device is a cudaGetSymbolAddress.
 * The caller still passes q / k / v / cu_seqlens / out (/ lse / bitmask).
 * Limit: a request whose tmx_attn_workspace_bytes() exceeds TMX_STATIC_SCRATCH_BYTES returns TMX_ERR_STATIC_SCRATCH
 * and launches nothing (every dense-causal layer and every TMX_KERNEL_TCGEN05 / _PAIR request fits at any length: their
 * scratch is the block index, 32 KB at 128K tokens; a TMX_KERNEL_AUTO triangle layer -- split-KV partials of the dense
 * tail rows -- fits up to about 16K tokens of Qwen3-8B shape, 130 MB at 128K does not).
 * Sharing: every call on a device uses the SAME array, which keeps the last call's contents; each call rewrites
 * what it reads (the block index is cleared by the call itself), so nothing has to be zeroed by the caller.  Calls
 * that overlap on the GPU (two streams, two threads) RACE on the array: serialise them, or use tmx_prefill_attn. */
#define TMX_STATIC_SCRATCH_BYTES (16u << 20)
int tmx_prefill_attn_static(void* out, const void* q, const void* k, const void* v,
                            const int32_t* cu_seqlens, int B, int total_T, int max_seqlen,
                            int Hq, int Hkv, int D, int64_t q_stride_t, int64_t kv_stride_t,
                            int64_t o_stride_t, float scale, int si, int sl, int last,
                            int mode, const uint32_t* bitmask, int64_t bitmask_stride_words,
                            float* lse, int kernel, tmx_stream_t stream);

/* ---- chunked prefill / prefix context / decode over the KV cache ----
 * The same attention when the call's query rows are only the TAIL of each sequence: sequence b
 * has seq_lens[b] keys (context + new tokens) and cu_seqlens_q[b+1] - cu_seqlens_q[b] query rows,
 * which are rows seq_lens[b] - q_len .. seq_lens[b] - 1 of the (seq_lens[b] x seq_lens[b]) mask
 * -- the triangle predicate's N is the full sequence length, as in the reference generator.
 * Replaces the reference backend's context path, flash_attn_varlen_func(q, key_cache, value_cache,
 * cu_seqlens_q, seqused_k, block_table, causal) (vllm/v1/attention/backends/flash_attn.py:518-548;
 * metadata fields block_table / seq_lens :120-126).  K/V addressing, exactly one of:
 *   block_table != NULL: k / v are the paged cache halves viewed as [num_blocks * page_size, Hkv*D]
 *     rows (FlashAttention's NHD layout, flash_attn.py:78-87; kv_stride_t = elements per slot,
 *     kv_rows = num_blocks * page_size); key j of sequence b sits in slot
 *     block_table[b * block_table_stride + j / page_size] * page_size + j % page_size.
 *     page_size: a multiple of 16 that divides 128, or a multiple of 128.  The tcgen05 kernel pulls
 *     each 128-key tile with one TMA box per page.  Cache memory behind the listed pages must hold
 *     finite values (vLLM zero-initialises it): masked keys still enter P x V with weight 0.
 *   cu_seqlens_k != NULL: k / v are contiguous [sum seq_lens, Hkv, D] buffers (kv_rows rows).
 * q_len == 1 rows are decode steps (served by the same kernels; correct, not decode-optimised).
 * mode: TMX_MODE_DENSE_CAUSAL or TMX_MODE_TRIANGLE.  workspace: tmx_attn_workspace_bytes() of
 * (B, total_Tq, max_query_len, ...) suffices.  lse (optional): float [Hq, total_Tq]. */
int tmx_chunked_prefill_attn(void* out, const void* q, const int32_t* cu_seqlens_q,
                             const int32_t* seq_lens, int B, int total_Tq, int max_query_len,
                             int Hq, int Hkv, int D, int64_t q_stride_t, int64_t o_stride_t,
                             const void* k, const void* v, int64_t kv_stride_t, int64_t kv_rows,
                             const int32_t* cu_seqlens_k, const int32_t* block_table,
                             int64_t block_table_stride, int page_size, float scale, int si,
                             int sl, int last, int mode, float* lse, void* workspace,
                             size_t workspace_bytes, int kernel, tmx_stream_t stream);

/* ---- the step in front of attention -------------------
 * One HBM pass over q/k(/v): per-head RMSNorm of q and k (weight [D], fp32 variance),
 * NeoX rotary embedding from cos_sin_cache[positions[t]] ([max_pos, D] bf16: cos | sin),
 * and, when k_cache/v_cache are given, the write of the new K (normed + rotated) and V
 * rows into the paged cache at slot_mapping[t] (slot < 0 = padding token, skipped).
 * Replaces Qwen3Attention.forward's q_norm / k_norm / rotary_emb calls
 * (vllm/model_executor/models/qwen3.py:166-174) and reshape_and_cache_flash
 * (vllm/v1/attention/backends/flash_attn.py:498-507; csrc/cache_kernels.cu) with the
 * rounding order of the reference's forward_native methods
 * (vllm/model_executor/layers/layernorm.py:170-206,
 *  vllm/model_executor/layers/rotary_embedding/base.py:66-95, common.py:29-47).
 * q [T, Hq*D], k/v [T, Hkv*D] bf16, token strides in elements; q_out/k_out may alias q/k.
 * Cache layout: element (block, offset, head, d) at
 *   block*cache_block_stride + offset*cache_token_stride + head*cache_head_stride + d
 * (covers FlashAttention's [2, num_blocks, block_size, Hkv, D] NHD layout,
 * flash_attn.py:78-87, and the HND permutation), bf16 cache only ("auto" kv_cache_dtype).
 * D must be 128; all strides multiples of 8 elements, pointers 16-byte aligned.
 * v / k_cache / v_cache / slot_mapping may all be null (no cache write).
 * positions[t] must lie in [0, max_pos): like the reference's CUDA rotary kernel
 * (csrc/pos_encoding_kernels.cu) the device code does not check it. */
int tmx_qk_norm_rope(void* q_out, void* k_out, const void* q, const void* k,
                     const void* q_weight, const void* k_weight, const int64_t* positions,
                     const void* cos_sin_cache, int max_pos, int T, int Hq, int Hkv, int D,
                     int64_t q_stride_t, int64_t k_stride_t, int64_t qo_stride_t,
                     int64_t ko_stride_t, float eps, const void* v, int64_t v_stride_t,
                     void* k_cache, void* v_cache, const int64_t* slot_mapping,
                     int64_t cache_block_stride, int64_t cache_token_stride,
                     int64_t cache_head_stride, int block_size, tmx_stream_t stream);

/* The cache write alone (key/value already final): ops.reshape_and_cache_flash
 * (vllm/_custom_ops.py reshape_and_cache_flash -> csrc/cache_kernels.cu), bf16 cache. */
int tmx_reshape_and_cache_flash(const void* key, const void* value, void* k_cache, void* v_cache,
                                const int64_t* slot_mapping, int T, int Hkv, int D,
                                int64_t k_stride_t, int64_t v_stride_t,
                                int64_t cache_block_stride, int64_t cache_token_stride,
                                int64_t cache_head_stride, int block_size, tmx_stream_t stream);

/* ---- the step after attention --------------------------------------------
 * o_proj as a row-parallel GEMM whose epilogue is the reduce-scatter: C = A W^T (bf16 in, fp32 accumulation in
 * TMEM, bf16 out) replaces RowParallelLinear.forward's matmul (vllm/model_executor/layers/linear.py:1197-1359;
 * call site `self.o_proj(attn_output)`, vllm/model_executor/models/qwen3.py:190), and instead of the
 * all-reduce at linear.py:1359 every output tile is written by TMA store into the staging buffer of the rank
 * that owns its token rows -- device pointers that may be PEER memory (NVLink / NVSwitch) -- then
 * tmx_sum_partials adds the n_slots partials of the owned rows.
 *   a [M, K] (row stride a_stride), w [N, K] (row stride w_stride): K contiguous; N % 256 == 0, K % 64 == 0.
 *   n_dest == 1: out_ptrs_host[0] is C [M, N] (row stride out_stride).
 *   n_dest  > 1: rows [d * rows_per_dest, (d+1) * rows_per_dest) of C go to out_ptrs_host[d], a contiguous
 *                [n_slots, rows_per_dest, N] bf16 buffer, at slot `slot` (rows_per_dest % 128 == 0).
 * out_ptrs_host is a HOST array of n_dest (<= 8) device pointers. */
int tmx_linear_bf16(void* const* out_ptrs_host, int n_dest, int slot, int n_slots, int64_t rows_per_dest,
                    const void* a, const void* w, int M, int N, int K, int64_t a_stride, int64_t w_stride,
                    int64_t out_stride, tmx_stream_t stream);
/* out[r, :] = sum_s partials[s, r, :]  (partials [n_slots, rows, N] bf16 contiguous, fp32 accumulation). */
int tmx_sum_partials(void* out, const void* partials, int n_slots, int64_t rows, int N, int64_t out_stride,
                     tmx_stream_t stream);

/* Number of kernels the last tmx_prefill_attn call on this thread enqueued
 * (bench.py's gpu_launches accounting). */
int tmx_last_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* TMX_H_ */

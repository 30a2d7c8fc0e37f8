// torch.library registration of the C ABI (include/tmx.h): the "thin torch-extension layer" of the boundary.
// Style of the reference's binding TU, csrc/torch_bindings.cpp:19,92-101 (ops.def("name(schema)") +
// ops.impl("name", torch::kCUDA, &fn)): each function only unpacks tensors, calls the extern "C" launcher on
// at::cuda::getCurrentCUDAStream() and turns a non-zero status into a c10::Error (Python RuntimeError).
// Registered ops are visible to torch.compile / CUDA-graph capture (the ctypes path in ops.py is opaque to both):
//   torch.ops.trianglemix.prefill_attn / chunked_prefill_attn / qk_norm_rope / reshape_and_cache_flash /
//   mask_bits / mask_additive / attn_workspace_bytes
// Built by csrc/Makefile target `torch` into vllm_trianglemix_b200/libtmx_torch.so (links libtmx.so, libtorch).
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/library.h>

#include "tmx.h"

namespace {

using at::Tensor;

void check(int status, const char* what) {
  TORCH_CHECK(status == TMX_OK, what, " failed: ", tmx_strerror(status),
              status == TMX_ERR_CUDA ? tmx_last_cuda_error() : "", " (status ", status, ")");
}
tmx_stream_t stream_of(const Tensor& t) {
  return reinterpret_cast<tmx_stream_t>(at::cuda::getCurrentCUDAStream(t.get_device()).stream());
}
void need_bf16_cuda(const Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor (there is no CPU path)");
  TORCH_CHECK(t.scalar_type() == at::kBFloat16, name, " must be bfloat16");
}
int dtype_code(const Tensor& t) {
  switch (t.scalar_type()) {
    case at::kFloat: return TMX_F32;
    case at::kBFloat16: return TMX_BF16;
    case at::kHalf: return TMX_F16;
    default: TORCH_CHECK(false, "mask dtype must be float32, bfloat16 or float16");
  }
  return -1;
}

int64_t attn_workspace_bytes(int64_t B, int64_t total_T, int64_t max_seqlen, int64_t Hq, int64_t Hkv, int64_t D,
                             int64_t si, int64_t sl, int64_t last, int64_t mode, int64_t kernel) {
  return (int64_t)tmx_attn_workspace_bytes((int)B, (int)total_T, (int)max_seqlen, (int)Hq, (int)Hkv, (int)D, (int)si,
                                           (int)sl, (int)last, (int)mode, (int)kernel);
}

// out [T,Hq,D], q [T,Hq,D], k/v [T,Hkv,D] bf16 (token-major, heads contiguous); cu_seqlens int32 [B+1];
// bitmask optional int32 [N, stride]; lse optional float [Hq, T]; workspace uint8.
void prefill_attn(Tensor& out, const Tensor& q, const Tensor& k, const Tensor& v, const Tensor& cu_seqlens,
                  int64_t max_seqlen, double scale, int64_t si, int64_t sl, int64_t last, int64_t mode,
                  const c10::optional<Tensor>& bitmask, const c10::optional<Tensor>& lse, Tensor& workspace,
                  int64_t kernel) {
  need_bf16_cuda(out, "out"); need_bf16_cuda(q, "q"); need_bf16_cuda(k, "k"); need_bf16_cuda(v, "v");
  TORCH_CHECK(q.dim() == 3 && k.dim() == 3 && v.sizes() == k.sizes() && out.sizes() == q.sizes(),
              "expected q/out [T,Hq,D], k/v [T,Hkv,D]");
  TORCH_CHECK(q.stride(2) == 1 && q.stride(1) == q.size(2) && k.stride(2) == 1 && k.stride(1) == k.size(2) &&
              v.stride(2) == 1 && v.stride(1) == v.size(2) && k.stride(0) == v.stride(0) && out.stride(2) == 1 &&
              out.stride(1) == out.size(2), "heads must be contiguous (token strides are free)");
  TORCH_CHECK(cu_seqlens.scalar_type() == at::kInt && cu_seqlens.is_contiguous() && cu_seqlens.is_cuda(),
              "cu_seqlens must be a contiguous int32 CUDA tensor");
  const c10::cuda::CUDAGuard guard(q.device());
  check(tmx_prefill_attn(out.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), cu_seqlens.data_ptr<int>(),
                         (int)cu_seqlens.numel() - 1, (int)q.size(0), (int)max_seqlen, (int)q.size(1), (int)k.size(1),
                         (int)q.size(2), q.stride(0), k.stride(0), out.stride(0), (float)scale, (int)si, (int)sl,
                         (int)last, (int)mode,
                         bitmask ? reinterpret_cast<const uint32_t*>(bitmask->data_ptr()) : nullptr,
                         bitmask ? bitmask->stride(0) : 0, lse ? lse->data_ptr<float>() : nullptr,
                         workspace.data_ptr(), (size_t)workspace.numel(), (int)kernel, stream_of(q)),
        "tmx_prefill_attn");
}

// Context path: k_cache / v_cache are the paged cache halves [num_blocks, block_size, Hkv, D] (NHD) with
// block_table int32 [B, max_blocks]; seq_lens int32 [B].
void chunked_prefill_attn(Tensor& out, const Tensor& q, const Tensor& cu_seqlens_q, const Tensor& seq_lens,
                          int64_t max_query_len, const Tensor& k_cache, const Tensor& v_cache,
                          const Tensor& block_table, double scale, int64_t si, int64_t sl, int64_t last, int64_t mode,
                          Tensor& workspace, int64_t kernel) {
  need_bf16_cuda(out, "out"); need_bf16_cuda(q, "q"); need_bf16_cuda(k_cache, "k_cache"); need_bf16_cuda(v_cache, "v_cache");
  TORCH_CHECK(k_cache.dim() == 4 && k_cache.sizes() == v_cache.sizes() && k_cache.strides() == v_cache.strides() &&
              k_cache.stride(3) == 1 && k_cache.stride(2) == k_cache.size(3) &&
              k_cache.stride(0) == k_cache.size(1) * k_cache.stride(1),
              "k_cache / v_cache must be NHD views [num_blocks, block_size, Hkv, D]");
  TORCH_CHECK(block_table.scalar_type() == at::kInt && block_table.dim() == 2 && block_table.stride(1) == 1,
              "block_table must be int32 [B, max_blocks]");
  const c10::cuda::CUDAGuard guard(q.device());
  check(tmx_chunked_prefill_attn(out.data_ptr(), q.data_ptr(), cu_seqlens_q.data_ptr<int>(), seq_lens.data_ptr<int>(),
                                 (int)seq_lens.numel(), (int)q.size(0), (int)max_query_len, (int)q.size(1),
                                 (int)k_cache.size(2), (int)q.size(2), q.stride(0), out.stride(0), k_cache.data_ptr(),
                                 v_cache.data_ptr(), k_cache.stride(1), k_cache.size(0) * k_cache.size(1), nullptr,
                                 block_table.data_ptr<int>(), block_table.stride(0), (int)k_cache.size(1), (float)scale,
                                 (int)si, (int)sl, (int)last, (int)mode, nullptr, workspace.data_ptr(),
                                 (size_t)workspace.numel(), (int)kernel, stream_of(q)),
        "tmx_chunked_prefill_attn");
}

// In place on q [T, Hq*D] / k [T, Hkv*D]; optional fused cache write (v, k_cache, v_cache, slot_mapping).
void qk_norm_rope(Tensor& q, Tensor& k, const Tensor& positions, const Tensor& q_weight, const Tensor& k_weight,
                  const Tensor& cos_sin_cache, double eps, int64_t head_dim, const c10::optional<Tensor>& v,
                  const c10::optional<Tensor>& k_cache, const c10::optional<Tensor>& v_cache,
                  const c10::optional<Tensor>& slot_mapping) {
  need_bf16_cuda(q, "q"); need_bf16_cuda(k, "k");
  TORCH_CHECK(positions.scalar_type() == at::kLong && positions.is_contiguous(), "positions must be contiguous int64");
  const bool cache = k_cache.has_value();
  TORCH_CHECK(!cache || (v && v_cache && slot_mapping), "the fused cache write needs v, k_cache, v_cache, slot_mapping");
  const c10::cuda::CUDAGuard guard(q.device());
  check(tmx_qk_norm_rope(q.data_ptr(), k.data_ptr(), q.data_ptr(), k.data_ptr(), q_weight.data_ptr(), k_weight.data_ptr(),
                         positions.data_ptr<int64_t>(), cos_sin_cache.data_ptr(), (int)cos_sin_cache.size(0),
                         (int)q.size(0), (int)(q.size(1) / head_dim), (int)(k.size(1) / head_dim), (int)head_dim,
                         q.stride(0), k.stride(0), q.stride(0), k.stride(0), (float)eps,
                         cache ? v->data_ptr() : nullptr, cache ? v->stride(0) : 0,
                         cache ? k_cache->data_ptr() : nullptr, cache ? v_cache->data_ptr() : nullptr,
                         cache ? slot_mapping->data_ptr<int64_t>() : nullptr, cache ? k_cache->stride(0) : 0,
                         cache ? k_cache->stride(1) : 0, cache ? k_cache->stride(2) : 0,
                         cache ? (int)k_cache->size(1) : 0, stream_of(q)),
        "tmx_qk_norm_rope");
}

void reshape_and_cache_flash(const Tensor& key, const Tensor& value, Tensor& k_cache, Tensor& v_cache,
                             const Tensor& slot_mapping) {
  need_bf16_cuda(key, "key"); need_bf16_cuda(value, "value");
  const c10::cuda::CUDAGuard guard(key.device());
  check(tmx_reshape_and_cache_flash(key.data_ptr(), value.data_ptr(), k_cache.data_ptr(), v_cache.data_ptr(),
                                    slot_mapping.data_ptr<int64_t>(), (int)std::min(slot_mapping.numel(), key.size(0)),
                                    (int)key.size(1), (int)key.size(2), key.stride(0), value.stride(0),
                                    k_cache.stride(0), k_cache.stride(1), k_cache.stride(2), (int)k_cache.size(1),
                                    stream_of(key)),
        "tmx_reshape_and_cache_flash");
}

void mask_bits(Tensor& out, int64_t n, int64_t si, int64_t sl, int64_t last, int64_t piece) {
  TORCH_CHECK(out.is_cuda() && out.scalar_type() == at::kInt && out.dim() == 2 && out.stride(1) == 1,
              "out must be an int32 CUDA tensor [N, stride]");
  const c10::cuda::CUDAGuard guard(out.device());
  check(tmx_mask_bits(reinterpret_cast<uint32_t*>(out.data_ptr()), out.stride(0), (int)n, (int)si, (int)sl, (int)last,
                      (int)piece, stream_of(out)), "tmx_mask_bits");
}

void mask_additive(Tensor& out, int64_t n, int64_t si, int64_t sl, int64_t last, int64_t mode) {
  TORCH_CHECK(out.is_cuda() && out.dim() == 2 && out.stride(1) == 1, "out must be a CUDA tensor [N, N]");
  const c10::cuda::CUDAGuard guard(out.device());
  check(tmx_mask_additive(out.data_ptr(), dtype_code(out), out.stride(0), (int)n, (int)si, (int)sl, (int)last,
                          (int)mode, stream_of(out)), "tmx_mask_additive");
}

}  // namespace

TORCH_LIBRARY(trianglemix, m) {
  m.def("attn_workspace_bytes(int B, int total_T, int max_seqlen, int Hq, int Hkv, int D, int si, int sl, int last, "
        "int mode, int kernel) -> int", &attn_workspace_bytes);
  m.def("prefill_attn(Tensor(a!) out, Tensor q, Tensor k, Tensor v, Tensor cu_seqlens, int max_seqlen, float scale, "
        "int si, int sl, int last, int mode, Tensor? bitmask, Tensor(b!)? lse, Tensor(c!) workspace, int kernel) -> ()");
  m.def("chunked_prefill_attn(Tensor(a!) out, Tensor q, Tensor cu_seqlens_q, Tensor seq_lens, int max_query_len, "
        "Tensor k_cache, Tensor v_cache, Tensor block_table, float scale, int si, int sl, int last, int mode, "
        "Tensor(b!) workspace, int kernel) -> ()");
  m.def("qk_norm_rope(Tensor(a!) q, Tensor(b!) k, Tensor positions, Tensor q_weight, Tensor k_weight, "
        "Tensor cos_sin_cache, float eps, int head_dim, Tensor? v, Tensor(c!)? k_cache, Tensor(d!)? v_cache, "
        "Tensor? slot_mapping) -> ()");
  m.def("reshape_and_cache_flash(Tensor key, Tensor value, Tensor(a!) k_cache, Tensor(b!) v_cache, "
        "Tensor slot_mapping) -> ()");
  m.def("mask_bits(Tensor(a!) out, int n, int si, int sl, int last, int piece) -> ()");
  m.def("mask_additive(Tensor(a!) out, int n, int si, int sl, int last, int mode) -> ()");
}

TORCH_LIBRARY_IMPL(trianglemix, CUDA, m) {
  m.impl("prefill_attn", &prefill_attn);
  m.impl("chunked_prefill_attn", &chunked_prefill_attn);
  m.impl("qk_norm_rope", &qk_norm_rope);
  m.impl("reshape_and_cache_flash", &reshape_and_cache_flash);
  m.impl("mask_bits", &mask_bits);
  m.impl("mask_additive", &mask_additive);
}

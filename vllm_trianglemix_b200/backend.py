"""vLLM-shaped attention backend for the TriangleMix prefill path.

Keeps the signatures of the reference's plug-in slot so that
``Attention(..., attn_backend=TriangleMixBackend, **extra_impl_args)``
(vllm/attention/layer.py:102,186-201) can install it the way
ChunkedLocalAttention / EncoderOnlyAttention install theirs:

* ``AttentionBackend`` statics          vllm/attention/backends/abstract.py:27-83
* ``AttentionImpl.__init__/forward``    vllm/attention/backends/abstract.py:142-171
* metadata fields                       vllm/v1/attention/backends/flash_attn.py:110-140
* in-place ``output`` contract          vllm/v1/attention/backends/flash_attn.py:449-458,549

The classes are duck-typed (no ``import vllm``: the vllm wheel in this image is an
unrelated upstream release).  The new K/V rows are written to the paged ``kv_cache`` at
``attn_metadata.slot_mapping`` (flash_attn.py:483-507).  A whole-sequence prefill attends the
call's own K/V (as ``_forward_encoder_attention`` does, flash_attn.py:578-633); a batch that
carries context keys (chunked prefill, prefix-cache hit, decode: ``seq_lens`` > query lengths)
reads all keys back from the cache through ``block_table`` (flash_attn.py:518-548).  TriangleMix parameters and ``layer_idx`` travel as extra impl kwargs,
the way ``dual_chunk_attention_config`` + ``layer_idx`` do in the reference's
vllm/model_executor/models/qwen3.py:150-153.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Tuple, Type

import torch

from . import ops
from ._lib import TmxError
from .trianglemix_attention import TriangleMixAttention, TriangleMixConfig


@dataclass
class TriangleMixMetadata:
    """Subset of FlashAttentionMetadata this path reads (flash_attn.py:110-140)."""
    num_actual_tokens: int
    max_query_len: int
    query_start_loc: torch.Tensor          # int32 [B+1] on the device
    max_seq_len: int
    seq_lens: Optional[torch.Tensor] = None          # int32 [B]: context + new tokens (flash_attn.py:124)
    causal: bool = True
    query_start_loc_cpu: Optional[torch.Tensor] = None
    slot_mapping: Optional[torch.Tensor] = None   # int64 [num_actual_tokens], flash_attn.py:127
    block_table: Optional[torch.Tensor] = None    # int32 [B, max_blocks], flash_attn.py:125
    seq_lens_cpu: Optional[torch.Tensor] = None
    # True when some sequence brings context keys (chunked prefill, prefix-cache hit, decode):
    # the keys then come from the paged cache.  None = unknown (decided from the CPU copies).
    has_context: Optional[bool] = None

    def context_present(self) -> bool:
        if self.has_context is not None:
            return bool(self.has_context)
        if self.max_seq_len != self.max_query_len:
            return True
        if self.seq_lens_cpu is not None and self.query_start_loc_cpu is not None:
            qlen = self.query_start_loc_cpu[1:] - self.query_start_loc_cpu[:-1]
            return bool((self.seq_lens_cpu[:qlen.numel()].to(qlen.dtype) != qlen).any())
        return False


class TriangleMixMetadataBuilder:
    """Builds TriangleMixMetadata from the fields of CommonAttentionMetadata
    (vllm/v1/attention/backends/utils.py:45-84), the way FlashAttentionMetadataBuilder.build does
    (vllm/v1/attention/backends/flash_attn.py:208-330)."""

    def __init__(self, kv_cache_spec=None, layer_names=None, vllm_config=None, device=None):
        self.kv_cache_spec = kv_cache_spec
        self.layer_names = layer_names
        self.vllm_config = vllm_config
        self.device = device

    def build(self, common_prefix_len: int = 0, common_attn_metadata=None, fast_build: bool = False):
        m = common_attn_metadata
        computed = getattr(m, "num_computed_tokens_cpu", None)
        has_ctx = None if computed is None else bool((computed[:int(getattr(m, "num_reqs", computed.numel()))] > 0).any())
        return TriangleMixMetadata(
            num_actual_tokens=int(m.num_actual_tokens), max_query_len=int(m.max_query_len),
            query_start_loc=m.query_start_loc, max_seq_len=int(getattr(m, "max_seq_len", m.max_query_len)),
            seq_lens=getattr(m, "seq_lens", None), causal=getattr(m, "causal", True),
            query_start_loc_cpu=getattr(m, "query_start_loc_cpu", None),
            slot_mapping=getattr(m, "slot_mapping", None),
            block_table=getattr(m, "block_table_tensor", getattr(m, "block_table", None)),
            seq_lens_cpu=getattr(m, "seq_lens_cpu", None), has_context=has_ctx)

    @staticmethod
    def from_seq_lens(seq_lens: List[int], device, query_lens: Optional[List[int]] = None,
                      block_table: Optional[torch.Tensor] = None,
                      slot_mapping: Optional[torch.Tensor] = None) -> TriangleMixMetadata:
        """Convenience for tests / benchmarks: a batch whose sequence b has seq_lens[b] keys, the last
        query_lens[b] of them being this call's query rows (default: all of them = whole-sequence prefill)."""
        qlens = list(seq_lens) if query_lens is None else list(query_lens)
        cu = [0]
        for n in qlens:
            cu.append(cu[-1] + int(n))
        cu_cpu = torch.tensor(cu, dtype=torch.int32)
        lens_cpu = torch.tensor(list(seq_lens), dtype=torch.int32)
        return TriangleMixMetadata(
            num_actual_tokens=cu[-1], max_query_len=max(qlens) if qlens else 0,
            query_start_loc=cu_cpu.to(device), max_seq_len=max(seq_lens) if seq_lens else 0,
            seq_lens=lens_cpu.to(device), causal=True, query_start_loc_cpu=cu_cpu,
            slot_mapping=slot_mapping, block_table=block_table, seq_lens_cpu=lens_cpu,
            has_context=any(int(a) != int(b) for a, b in zip(seq_lens, qlens)))


class TriangleMixImpl:
    """AttentionImpl for TriangleMix prefill (signature: abstract.py:142-171)."""

    def __init__(
        self,
        num_heads: int,
        head_size: int,
        scale: float,
        num_kv_heads: Optional[int] = None,
        alibi_slopes: Optional[List[float]] = None,
        sliding_window: Optional[int] = None,
        kv_cache_dtype: str = "auto",
        logits_soft_cap: Optional[float] = None,
        attn_type: str = "decoder",
        kv_sharing_target_layer_name: Optional[str] = None,
        # ---- extra_impl_args (layer.py:186-201) ----
        trianglemix_config: Optional[TriangleMixConfig] = None,
        layer_idx: Optional[int] = None,
        num_layers: Optional[int] = None,
        kernel: str = "auto",
    ) -> None:
        if alibi_slopes is not None or logits_soft_cap:
            raise NotImplementedError("TriangleMixImpl supports neither ALiBi nor logits soft-capping")
        if sliding_window is not None:
            raise NotImplementedError("model-level sliding_window is not combined with TriangleMix; "
                                      "set TriangleMixConfig.sliding_window_size instead")
        if attn_type != "decoder":
            raise NotImplementedError("TriangleMixImpl is decoder self-attention only")
        if kv_sharing_target_layer_name is not None:
            raise NotImplementedError("KV sharing is not supported")
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        self.num_queries_per_kv = self.num_heads // self.num_kv_heads
        self.kv_cache_dtype = kv_cache_dtype
        self.layer_idx = layer_idx or 0            # reference ascend_trianglemix.py:43
        self.kernel = kernel
        self.trianglemix_attn = None
        if trianglemix_config is not None and num_layers is not None:
            self.trianglemix_attn = TriangleMixAttention(trianglemix_config, num_layers)

    def mask_mode(self) -> Tuple[int, int, int, int]:
        """(mode, si, sl, last): the per-layer dense-vs-triangle switch
        (reference trianglemix_attention.py:263-278) feeding the kernel's mask_mode.
        Without a helper the layer is plain causal attention (what the stock backend
        the inert hook falls through to computes, reference qwen3.py:185-189)."""
        if self.trianglemix_attn is None:
            return (ops.DENSE_CAUSAL, 0, 0, 0)
        si, sl, last = self.trianglemix_attn.config.mask_params()
        return (self.trianglemix_attn.mask_mode(self.layer_idx), si, sl, last)

    def forward(
        self,
        layer,
        query: torch.Tensor,
        key: torch.Tensor,
        value: torch.Tensor,
        kv_cache: Optional[torch.Tensor],
        attn_metadata: Optional[TriangleMixMetadata],
        output: Optional[torch.Tensor] = None,
        output_scale: Optional[torch.Tensor] = None,
        output_block_scale: Optional[torch.Tensor] = None,
    ) -> torch.Tensor:
        assert output is not None, "Output tensor must be provided."           # flash_attn.py:449
        if output_scale is not None or output_block_scale is not None:        # flash_attn.py:451-454
            raise NotImplementedError("fused output quantization is not supported for TriangleMixImpl")
        if attn_metadata is None:                                              # profiling run, :456-458
            return output
        if not attn_metadata.causal:
            raise NotImplementedError("TriangleMixImpl is causal prefill only")
        t = attn_metadata.num_actual_tokens
        q = query[:t].view(t, self.num_heads, self.head_size)
        k = key[:t].view(t, self.num_kv_heads, self.head_size)
        v = value[:t].view(t, self.num_kv_heads, self.head_size)
        out = output[:t].view(t, self.num_heads, self.head_size)
        slot_mapping = getattr(attn_metadata, "slot_mapping", None)
        if kv_cache is not None and kv_cache.numel() > 0 and slot_mapping is not None:
            # store the new K/V rows in the paged cache (flash_attn.py:483-507)
            key_cache, value_cache = kv_cache.unbind(0)
            ops.reshape_and_cache_flash(k, v, key_cache, value_cache, slot_mapping, self.kv_cache_dtype)
        return self.attend(q, k, v, kv_cache, attn_metadata, out, output)

    def attend(self, q, k, v, kv_cache, attn_metadata: TriangleMixMetadata, out, output):
        """The attention itself, after the new K/V rows are in the cache (callers that fuse the
        cache write into an earlier kernel -- Qwen3TriangleMixAttention -- enter here).
        q/k/v/out: [T, H, D] views of the actual tokens; returns `output`."""
        mode, si, sl, last = self.mask_mode()
        cu_cpu = attn_metadata.query_start_loc_cpu
        if cu_cpu is not None and cu_cpu.numel() > 1:
            # the kernels size their grids from max_query_len: a value below the true maximum would leave query
            # tiles without a worker (checked here from the CPU copy, no device sync)
            true_max = int((cu_cpu[1:] - cu_cpu[:-1]).max())
            if attn_metadata.max_query_len < true_max:
                raise ValueError("attn_metadata.max_query_len (%d) is smaller than the longest query (%d)" %
                                 (attn_metadata.max_query_len, true_max))
        if attn_metadata.context_present():
            # Chunked prefill / prefix-cache hit / decode: the keys are the cache's, through the block
            # table (flash_attn.py:518-548).  Query rows are the LAST rows of each sequence, so the
            # triangle predicate sees the full sequence length.
            if kv_cache is None or kv_cache.numel() == 0 or attn_metadata.block_table is None \
                    or attn_metadata.seq_lens is None:
                raise NotImplementedError(
                    "this batch carries context keys (seq_lens != query lengths: chunked prefill, a prefix-cache "
                    "hit or a decode step) but no kv_cache / block_table / seq_lens to read them from")
            key_cache, value_cache = kv_cache.unbind(0)
            self._context_prefill(q, out, key_cache, value_cache, attn_metadata, mode, si, sl, last)
            return output
        self._whole_sequence_prefill(q, k, v, out, attn_metadata, mode, si, sl, last)
        return output

    # The reference helper hands out no mask at seq_len <= 2048 (trianglemix_attention.py:330-332): such a
    # sequence runs on the stock causal backend even in a triangle layer.
    SHORT_SEQ_DENSE = 2048

    def _context_prefill(self, q, out, key_cache, value_cache, md: TriangleMixMetadata, mode, si, sl, last) -> None:
        """Keys from the paged cache.  In a triangle layer a sequence of <= 2048 tokens (so far) is dense-causal
        (the reference's threshold); runs of consecutive sequences of one class go out as one kernel call."""
        def call(b0, b1, t0, t1, qmax, m):
            whole = b0 == 0 and b1 == md.seq_lens.numel()
            cu = md.query_start_loc if whole else (md.query_start_loc[b0:b1 + 1] - md.query_start_loc[b0]).contiguous()
            ops.chunked_prefill_attn(q[t0:t1], cu, md.seq_lens[b0:b1], qmax, self.scale, si, sl, last, m,
                                     key_cache=key_cache, value_cache=value_cache, block_table=md.block_table[b0:b1],
                                     out=out[t0:t1], kernel=self.kernel)

        nseq = md.seq_lens.numel()
        if mode != ops.TRIANGLE or md.max_seq_len <= self.SHORT_SEQ_DENSE:
            call(0, nseq, 0, md.num_actual_tokens, md.max_query_len, ops.DENSE_CAUSAL if mode == ops.TRIANGLE else mode)
            return
        if md.seq_lens_cpu is None or md.query_start_loc_cpu is None:
            raise NotImplementedError("a triangle layer needs the CPU copies of seq_lens / query_start_loc to separate "
                                      "sequences <= 2048 tokens (dense by the reference's threshold) from longer ones")
        lens = md.seq_lens_cpu[:nseq].tolist()
        starts = md.query_start_loc_cpu[:nseq + 1].tolist()
        b0 = 0
        while b0 < nseq:
            short = lens[b0] <= self.SHORT_SEQ_DENSE
            b1 = b0
            while b1 < nseq and (lens[b1] <= self.SHORT_SEQ_DENSE) == short:
                b1 += 1
            if starts[b1] > starts[b0]:
                qmax = max(starts[b + 1] - starts[b] for b in range(b0, b1))
                call(b0, b1, starts[b0], starts[b1], qmax, ops.DENSE_CAUSAL if short else ops.TRIANGLE)
            b0 = b1

    def _whole_sequence_prefill(self, q, k, v, out, md: TriangleMixMetadata, mode, si, sl, last) -> None:
        cu, cu_cpu = md.query_start_loc, md.query_start_loc_cpu
        if mode != ops.TRIANGLE or cu_cpu is None:
            if mode == ops.TRIANGLE and md.max_query_len <= self.SHORT_SEQ_DENSE:
                mode = ops.DENSE_CAUSAL
            ops.prefill_attn(q, k, v, cu, md.max_query_len, self.scale, si, sl, last, mode, out=out,
                             kernel=self.kernel)
            return
        # triangle layer: sequences <= 2048 tokens are dense-causal, longer ones triangle.  Runs of
        # consecutive sequences of one class are contiguous token ranges -> one kernel call per run.
        lens = (cu_cpu[1:] - cu_cpu[:-1]).tolist()
        starts = cu_cpu.tolist()
        b0 = 0
        while b0 < len(lens):
            short = lens[b0] <= self.SHORT_SEQ_DENSE
            b1 = b0
            while b1 < len(lens) and (lens[b1] <= self.SHORT_SEQ_DENSE) == short:
                b1 += 1
            t0, t1 = starts[b0], starts[b1]
            if t1 > t0:
                sub_cu = cu if (b0 == 0 and b1 == len(lens)) else (cu[b0:b1 + 1] - cu[b0]).contiguous()
                ops.prefill_attn(q[t0:t1], k[t0:t1], v[t0:t1], sub_cu, max(lens[b0:b1]), self.scale, si, sl, last,
                                 ops.DENSE_CAUSAL if short else ops.TRIANGLE, out=out[t0:t1], kernel=self.kernel)
            b0 = b1


class TriangleMixBackend:
    """AttentionBackend statics (abstract.py:27-83; FlashAttention's values flash_attn.py:37-88)."""

    accept_output_buffer: bool = True
    supports_quant_query_input: bool = False

    @staticmethod
    def get_name() -> str:
        return "TRIANGLEMIX_B200"

    @staticmethod
    def get_impl_cls() -> Type[TriangleMixImpl]:
        return TriangleMixImpl

    @staticmethod
    def get_metadata_cls() -> Type[TriangleMixMetadata]:
        return TriangleMixMetadata

    @classmethod
    def make_metadata(cls, *args, **kwargs) -> TriangleMixMetadata:
        return cls.get_metadata_cls()(*args, **kwargs)

    @staticmethod
    def get_builder_cls() -> Type[TriangleMixMetadataBuilder]:
        return TriangleMixMetadataBuilder

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int,
                           cache_dtype_str: str = "auto") -> Tuple[int, ...]:
        if block_size % 16 != 0:
            raise ValueError("Block size must be a multiple of 16.")
        return (2, num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return [128]

    @staticmethod
    def get_supported_dtypes() -> List[torch.dtype]:
        return [torch.bfloat16]

    @classmethod
    def full_cls_name(cls) -> Tuple[str, str]:
        return (cls.__module__, cls.__qualname__)

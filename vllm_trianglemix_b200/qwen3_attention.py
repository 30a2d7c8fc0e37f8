"""The attention block of the reference's Qwen3 layer between ``qkv_proj`` and ``o_proj``,
with the TriangleMix hook actually wired.

Reference: vllm/model_executor/models/qwen3.py
  :62-114   constructor arguments (heads per tensor-parallel rank :85-98, scaling :102,
            ``trianglemix_config / layer_idx / num_layers`` :79-81,105-114)
  :160-190  forward: split -> q_norm / k_norm per head -> rotary_emb -> [mask hook] -> attn
What differs from the reference, on purpose (its hook is inert):
  * the per-layer dense-vs-triangle decision reaches the attention kernel (the reference
    computes a mask at :177-184 and drops it);
  * the sequence length comes from ``attn_metadata.query_start_loc`` instead of ``q.shape[1]``
    (:178 reads the hidden size of vLLM's [num_tokens, q_size] layout);
  * q_norm + k_norm + rotary_emb (+ the KV-cache write of flash_attn.py:498-507) are ONE
    kernel (``tmx_qk_norm_rope``) working in place on the packed qkv buffer (as the reference's
    CUDA rotary_embedding does), so q/k/v stay views of one buffer with one token stride.
The two GEMMs either side (``qkv_proj``, ``o_proj``) are library GEMMs and stay with the caller.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from . import ops
from .backend import TriangleMixImpl, TriangleMixMetadata
from .trianglemix_attention import TriangleMixConfig


def compute_cos_sin_cache(head_size: int, max_position: int, base: float, dtype=torch.bfloat16,
                          device=None) -> torch.Tensor:
    """[max_position, head_size] cos | sin table -- RotaryEmbedding._compute_cos_sin_cache
    (vllm/model_executor/layers/rotary_embedding/base.py:42-58), built once at model
    construction in fp32 on the host and cast, like the reference's registered buffer (:36-40)."""
    inv_freq = 1.0 / (base ** (torch.arange(0, head_size, 2, dtype=torch.float) / head_size))
    t = torch.arange(max_position, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    cache = torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype)
    return cache if device is None else cache.to(device)


class Qwen3TriangleMixAttention(nn.Module):
    """Drop-in for the body of ``Qwen3Attention.forward`` between the two projections."""

    def __init__(self, num_heads: int, num_kv_heads: int, head_dim: int = 128, max_position: int = 4096 * 32,
                 rms_norm_eps: float = 1e-06, rope_theta: float = 10000,
                 trianglemix_config: Optional[TriangleMixConfig] = None, layer_idx: Optional[int] = None,
                 num_layers: Optional[int] = None, tp_size: int = 1, device=None, kernel: str = "auto") -> None:
        super().__init__()
        # heads per rank exactly as qwen3.py:85-98
        assert num_heads % tp_size == 0
        if num_kv_heads >= tp_size:
            assert num_kv_heads % tp_size == 0
        else:
            assert tp_size % num_kv_heads == 0
        self.num_heads = num_heads // tp_size
        self.num_kv_heads = max(1, num_kv_heads // tp_size)
        self.head_dim = head_dim
        self.q_size = self.num_heads * head_dim
        self.kv_size = self.num_kv_heads * head_dim
        self.scaling = head_dim ** -0.5                                        # qwen3.py:102
        self.rms_norm_eps = rms_norm_eps
        self.layer_idx = layer_idx or 0
        self.q_norm_weight = nn.Parameter(torch.ones(head_dim, dtype=torch.bfloat16, device=device),
                                          requires_grad=False)                 # RMSNorm weight, layernorm.py:150-153
        self.k_norm_weight = nn.Parameter(torch.ones(head_dim, dtype=torch.bfloat16, device=device),
                                          requires_grad=False)
        self.register_buffer("cos_sin_cache",
                             compute_cos_sin_cache(head_dim, max_position, rope_theta, device=device),
                             persistent=False)
        self.impl = TriangleMixImpl(self.num_heads, head_dim, self.scaling, num_kv_heads=self.num_kv_heads,
                                    trianglemix_config=trianglemix_config, layer_idx=layer_idx,
                                    num_layers=num_layers, kernel=kernel)

    def forward(self, positions: torch.Tensor, qkv: torch.Tensor, kv_cache: Optional[torch.Tensor] = None,
                attn_metadata: Optional[TriangleMixMetadata] = None,
                output: Optional[torch.Tensor] = None) -> torch.Tensor:
        """qkv: [T, q_size + 2*kv_size] bf16 (the output of ``qkv_proj``); returns attn_output
        [T, q_size] (the input of ``o_proj``)."""
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)  # views, qwen3.py:161
        if output is None:
            output = torch.empty((qkv.shape[0], self.q_size), dtype=qkv.dtype, device=qkv.device)
        cache_args = {}
        slot_mapping = getattr(attn_metadata, "slot_mapping", None)
        fused_cache = kv_cache is not None and kv_cache.numel() > 0 and slot_mapping is not None
        if fused_cache:
            key_cache, value_cache = kv_cache.unbind(0)                         # flash_attn.py:483
            cache_args = dict(v=v, key_cache=key_cache, value_cache=value_cache, slot_mapping=slot_mapping)
        q, k = ops.qk_norm_rope(positions, q, k, self.q_norm_weight, self.k_norm_weight, self.cos_sin_cache,
                                self.rms_norm_eps, self.head_dim, inplace=True, **cache_args)
        if not fused_cache or attn_metadata is None:
            return self.impl.forward(None, q, k, v, kv_cache, attn_metadata, output=output)
        # the cache rows are already written by the fused kernel: enter the impl after its cache write
        t, impl = attn_metadata.num_actual_tokens, self.impl
        return impl.attend(q[:t].view(t, impl.num_heads, impl.head_size), k[:t].view(t, impl.num_kv_heads, impl.head_size),
                           v[:t].view(t, impl.num_kv_heads, impl.head_size), kv_cache, attn_metadata,
                           output[:t].view(t, impl.num_heads, impl.head_size), output)

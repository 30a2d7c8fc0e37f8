"""Drop-in for the reference's vllm/attention/ascend_trianglemix.py.

`AscendTriangleMixAttention.forward` keeps the reference's signature and its
observable behaviour, including the quirks:
  * the helper returns no mask at seq_len <= 2048, and with no caller mask the
    op is then plain BIDIRECTIONAL attention (reference :96-97);
  * a caller `attn_mask` is combined with the layer's mask (reference :76-84).
What changes is how it is computed: one flash-style CUDA kernel that skips every
fully-masked tile and never materialises (B,H,S,S) scores or an (S,S) mask.
"""
from __future__ import annotations

import warnings
from typing import Optional

import torch
from torch import nn

from . import ops
from ._lib import TmxError
from .trianglemix_attention import (TriangleMixAttention, TriangleMixConfig,  # noqa: F401
                                    TriangleMixMaskGenerator)


class AscendTriangleMixAttention(nn.Module):
    """Mirror of reference ascend_trianglemix.py:16-108.

    Additive extension: ``num_kv_heads`` (GQA; the reference op is MHA-only and
    raises on mismatched head counts, reference :63-65,93)."""

    def __init__(
        self,
        num_heads: int,
        head_dim: int,
        scale: float,
        trianglemix_config: Optional[TriangleMixConfig] = None,
        layer_idx: Optional[int] = None,
        num_layers: Optional[int] = None,
        num_kv_heads: Optional[int] = None,
        kernel: str = "auto",
    ):
        super().__init__()
        self.num_heads = num_heads
        self.head_dim = head_dim
        self.scale = scale
        self.layer_idx = layer_idx or 0
        self.num_kv_heads = num_kv_heads
        self.kernel = kernel
        self.trianglemix_attn = None
        if trianglemix_config is not None and num_layers is not None:
            self.trianglemix_attn = TriangleMixAttention(config=trianglemix_config, num_layers=num_layers)

    def _layer_mask(self, seq_len: int):
        """(mode, si, sl, last) of the mask the helper would return, or None (reference :330-332)."""
        helper = self.trianglemix_attn
        if helper is None or seq_len <= 2048:
            return None
        si, sl, last = helper.config.mask_params()
        return (helper.mask_mode(self.layer_idx), si, sl, last)

    def forward(self, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                attn_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        if query.dim() != 4:
            raise TmxError("query must be (batch, seq_len, num_heads, head_dim), got %r" % (tuple(query.shape),))
        b, s, hq, d = query.shape
        hkv = key.shape[2]
        if self.num_kv_heads is None and hkv != hq:
            # the reference fails here with a shape error from torch.matmul (reference :93)
            raise TmxError("key has %d heads but query has %d; pass num_kv_heads= to enable GQA" % (hkv, hq))
        in_dtype = query.dtype
        if in_dtype != torch.bfloat16:
            warnings.warn("AscendTriangleMixAttention computes in bfloat16 on B200; %s inputs are cast "
                          "and the result is cast back" % in_dtype, stacklevel=2)
        q = query.to(torch.bfloat16).reshape(b * s, hq, d)
        k = key.to(torch.bfloat16).reshape(b * s, hkv, d)
        v = value.to(torch.bfloat16).reshape(b * s, hkv, d)
        cu = torch.arange(0, (b + 1) * s, s, dtype=torch.int32, device=query.device)

        layer = self._layer_mask(s)
        bitmask = None
        mode, si, sl, last = ops.NONE, 0, 0, 0
        if layer is not None:
            mode, si, sl, last = layer
        if attn_mask is not None:
            spec = getattr(attn_mask, "_tmx_mask_spec", None)
            if spec is not None and spec[5] != attn_mask._version:
                spec = None                            # edited in place since the generator made it
            if spec is not None and spec[0] == s and (
                    layer is None or (spec[4], spec[1], spec[2], spec[3]) == tuple(layer)):
                _, si, sl, last, mode, _ = spec        # a mask made by our generator: analytic path
            else:
                bitmask = self._bits_of_user_mask(attn_mask, s)
                if layer is not None:                  # attn_mask + tri_mask  ==  AND of the bit sets
                    bitmask = bitmask & ops.mask_bits(s, si, sl, last,
                                                      "triangle" if mode == ops.TRIANGLE else "dense",
                                                      query.device)
                    bitmask = bitmask.contiguous()
        out = ops.prefill_attn(q, k, v, cu, s, self.scale, si, sl, last, mode, bitmask=bitmask,
                               kernel=self.kernel)
        out = out.view(b, s, hq, d)
        return out if in_dtype == torch.bfloat16 else out.to(in_dtype)

    @staticmethod
    def _bits_of_user_mask(attn_mask: torch.Tensor, s: int) -> torch.Tensor:
        if attn_mask.dim() != 2 or tuple(attn_mask.shape) != (s, s):
            raise TmxError("attn_mask must be an additive (seq_len, seq_len) tensor of 0 / -inf, got %r"
                           % (tuple(attn_mask.shape),))
        if attn_mask.dtype not in (torch.float32, torch.bfloat16, torch.float16):
            raise TmxError("attn_mask dtype must be float32/bfloat16/float16, got %s" % attn_mask.dtype)
        bits, bad = ops.bits_from_additive(attn_mask)
        if int(bad.item()) != 0:
            raise TmxError("attn_mask holds %d values other than 0 and -inf (or <= -1e4); only such additive "
                           "masks are supported (no eager fallback)" % int(bad.item()))
        return bits


class AscendNPUTriangleMixOptimizer:
    """Mirror of reference ascend_trianglemix.py:111-190 (mask rewrite only; the reference's
    `sparse_attention_kernel` is broken as written -- SURVEY 0.2 -- and is not reproduced)."""

    @staticmethod
    def optimize_mask_for_npu(mask: torch.Tensor, block_size: int = 64) -> torch.Tensor:
        if mask.numel() == 0:
            return mask
        return ops.rewrite_neg_inf(mask, -1e6)


def create_trianglemix_config_from_hf(hf_config, use_trianglemix: bool = False,
                                      num_triangle_layers: Optional[int] = None) -> Optional[TriangleMixConfig]:
    """Mirror of reference ascend_trianglemix.py:193-224 (last = max(64, 2 * sliding_window))."""
    if not use_trianglemix:
        return None
    num_sink_tokens = getattr(hf_config, "num_sink_tokens", 4)
    sliding_window_size = getattr(hf_config, "sliding_window", 32)
    num_last_tokens = max(64, sliding_window_size * 2)
    return TriangleMixConfig(
        num_sink_tokens=num_sink_tokens,
        sliding_window_size=sliding_window_size,
        num_last_tokens=num_last_tokens,
        num_triangle_layers=num_triangle_layers,
    )

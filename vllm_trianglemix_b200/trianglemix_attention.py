"""Drop-in for the reference's vllm/attention/trianglemix_attention.py.

Same class names, constructor arguments, method names, cache key and return
values; the masks themselves come from CUDA kernels (csrc/tmx_mask.cu) instead
of 2*N^2 Python iterations.  Native additions (not in the reference): packed
bitmask and block-index accessors, which are what the attention kernel consumes
-- the N x N tensors exist only for API compatibility.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import ops


class TriangleMixConfig:
    """Mirror of reference trianglemix_attention.py:11-36 (field for field)."""

    def __init__(
        self,
        num_sink_tokens: int = 4,
        sliding_window_size: int = 32,
        num_last_tokens: int = 64,
        num_triangle_layers: Optional[int] = None,
        triangle_layer_indices: Optional[list[int]] = None,
    ):
        self.num_sink_tokens = num_sink_tokens
        self.sliding_window_size = sliding_window_size
        self.num_last_tokens = num_last_tokens
        self.num_triangle_layers = num_triangle_layers
        self.triangle_layer_indices = triangle_layer_indices or []

    def mask_params(self):
        return (int(self.num_sink_tokens), int(self.sliding_window_size), int(self.num_last_tokens))


class TriangleMixMaskGenerator(nn.Module):
    """Mirror of reference trianglemix_attention.py:39-282."""

    def __init__(self, config: TriangleMixConfig):
        super().__init__()
        self.config = config
        self._mask_cache = {}

    # -- the five 0/1 pieces (reference :53, :93, :133, :173, :200) ------------
    def _piece(self, piece: str, seq_len: int, device, dtype) -> torch.Tensor:
        si, sl, last = self.config.mask_params()
        return ops.mask_dense01(seq_len, si, sl, last, piece, device, dtype)

    def _create_streaming_mask(self, seq_len: int, device: torch.device, dtype: torch.dtype) -> torch.Tensor:
        return self._piece("streaming", seq_len, device, dtype)

    def _create_last_mask(self, seq_len: int, device: torch.device, dtype: torch.dtype) -> torch.Tensor:
        return self._piece("last", seq_len, device, dtype)

    def _create_middle_mask(self, seq_len: int, device: torch.device, dtype: torch.dtype) -> torch.Tensor:
        return self._piece("middle", seq_len, device, dtype)

    def _create_triangle_mask(self, seq_len: int, device: torch.device, dtype: torch.dtype) -> torch.Tensor:
        return self._piece("triangle", seq_len, device, dtype)

    def _create_dense_mask(self, seq_len: int, device: torch.device, dtype: torch.dtype) -> torch.Tensor:
        return self._piece("dense", seq_len, device, dtype)

    # -- additive mask with the reference's cache (reference :221-261) ----------------
    def get_attention_mask(self, seq_len: int, layer_idx: int, device: torch.device,
                           dtype: torch.dtype, use_triangle: bool = False) -> torch.Tensor:
        cache_key = (seq_len, layer_idx, str(device), str(dtype), use_triangle)
        hit = self._mask_cache.get(cache_key)
        if hit is not None:
            return hit
        si, sl, last = self.config.mask_params()
        mask = ops.mask_additive(seq_len, si, sl, last,
                                 ops.TRIANGLE if use_triangle else ops.DENSE_CAUSAL, device, dtype)
        # tag the tensor so the attention op can recognise a generator-made mask and use the
        # block index instead of reading N x N values back
        # (the tag carries the tensor's version counter: an in-place edit by the caller voids it)
        mask._tmx_mask_spec = (seq_len, si, sl, last, ops.TRIANGLE if use_triangle else ops.DENSE_CAUSAL,
                               mask._version)
        self._mask_cache[cache_key] = mask
        return mask

    def should_use_triangle(self, layer_idx: int) -> bool:
        """Reference :263-278: explicit indices win, else first-K, else dense."""
        if self.config.triangle_layer_indices:
            return layer_idx in self.config.triangle_layer_indices
        elif self.config.num_triangle_layers is not None:
            return layer_idx < self.config.num_triangle_layers
        return False

    def clear_cache(self):
        self._mask_cache.clear()

    # -- native forms (additions) -----------------------------------------------------
    def mask_mode(self, layer_idx: int) -> int:
        return ops.TRIANGLE if self.should_use_triangle(layer_idx) else ops.DENSE_CAUSAL

    def get_bitmask(self, seq_len: int, device, use_triangle: bool = True, piece: Optional[str] = None):
        """Packed int32 bitmask [N, stride]; N^2/8 bytes instead of N^2*sizeof(dtype)."""
        si, sl, last = self.config.mask_params()
        piece = piece or ("triangle" if use_triangle else "dense")
        return ops.mask_bits(seq_len, si, sl, last, piece, device)

    def get_block_mask(self, cu_seqlens: torch.Tensor, total_tokens: int, use_triangle: bool = True,
                       block_m: int = 64, block_n: int = 64):
        """Block index: per query tile at most two KV tile ranges (O(N/block_m) work)."""
        si, sl, last = self.config.mask_params()
        return ops.block_index(cu_seqlens, total_tokens, block_m, block_n, si, sl, last,
                               ops.TRIANGLE if use_triangle else ops.DENSE_CAUSAL)


class TriangleMixAttention:
    """Mirror of reference trianglemix_attention.py:285-340."""

    def __init__(self, config: TriangleMixConfig, num_layers: int):
        self.config = config
        self.num_layers = num_layers
        self.mask_generator = TriangleMixMaskGenerator(config)
        # reference :305-307 -- the first-K default is written back into the shared config
        if config.num_triangle_layers is not None and not config.triangle_layer_indices:
            self.config.triangle_layer_indices = list(range(config.num_triangle_layers))

    def get_attn_mask(self, seq_len: int, layer_idx: int, device: torch.device,
                      dtype: torch.dtype = torch.float32) -> Optional[torch.Tensor]:
        use_triangle = self.mask_generator.should_use_triangle(layer_idx)
        if seq_len <= 2048:          # reference :330-332
            return None
        return self.mask_generator.get_attention_mask(
            seq_len=seq_len, layer_idx=layer_idx, device=device, dtype=dtype, use_triangle=use_triangle)

    # native addition: what the kernel needs instead of an N x N tensor
    def mask_mode(self, layer_idx: int) -> int:
        return self.mask_generator.mask_mode(layer_idx)


class TriangleMixAnalyzer(nn.Module):
    """Mirror of reference trianglemix_attention.py:344-409 (plain Python, no kernel)."""

    def __init__(self, num_layers: int):
        super().__init__()
        self.num_layers = num_layers
        self.middle_qk_gradients = [[] for _ in range(num_layers)]

    def record_middle_qk_gradient(self, layer_idx: int, gradient: torch.Tensor):
        if layer_idx < self.num_layers:
            self.middle_qk_gradients[layer_idx].append(gradient.detach().float().mean().item())

    def get_triangle_layers(self, num_triangle_layers: int) -> list[int]:
        means = []
        for layer_idx, samples in enumerate(self.middle_qk_gradients):
            means.append((sum(samples) / len(samples)) if samples else float("inf"))
        order = sorted(range(self.num_layers), key=lambda i: means[i])   # stable: ties by index
        return sorted(order[:num_triangle_layers])

    def reset(self):
        self.middle_qk_gradients = [[] for _ in range(self.num_layers)]
